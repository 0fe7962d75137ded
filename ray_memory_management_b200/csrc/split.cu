// Streaming split: the map side of the all-to-all exchange as one pass over the input.
//
// Replaces the M x N map-output objects of the reference's shuffle and their pull by the
// reducers (python/ray/data/_internal/shuffle.py:81-112, _internal/arrow_block.py:387-393;
// src/ray/object_manager/object_manager.cc:336-559): the rows of a block go, in input order,
// straight into the receive buffer of the GPU that owns their key range -- no partitioned copy
// of the block, no permutation, no random reads.
//
//   ds_split_plan   reads the key column once: destination of every row (one byte, binary
//                   search over the splitters in shared memory, null keys -> last partition)
//                   and, per tile of DS_SPLIT_TILE_ROWS rows and per destination, how many of
//                   the destination's rows lie in earlier tiles (exclusive scan, in place).
//   ds_split_rows   one launch per column.  A persistent CTA per SM walks tiles of T rows:
//                   the tile arrives with ONE bulk copy (cp.async.bulk global -> shared,
//                   mbarrier tracked) while the CTA ranks the rows stably per destination
//                   (warp ballots); one thread per row then moves the row inside shared memory
//                   to its place in its destination's span (laid out with the byte alignment
//                   of the span's global address), and every span leaves with one bulk store
//                   (cp.async.bulk shared -> global) to dst_bases[d] + offset -- peer-mapped
//                   memory of another GPU over NVLink, or this GPU's own buffer.  A split is a
//                   radix pass on whole rows with digit = destination; it is stable by
//                   construction, which is what makes the global result the stable sort.
//
// Algorithmic bytes per row: plan K (key) + 1; split 1 + 2 * W (W over NVLink for remote rows).
#include "ds_async.cuh"
#include "ds_common.cuh"
#include "ds_keys.cuh"
#include "gather_bulk.cuh"

namespace ds {

constexpr int SP_THREADS = 256;
constexpr int SP_TILE0 = DS_SPLIT_TILE_ROWS;  // granularity of the tile offsets
constexpr int SP_MAX_PARTS = DS_MAX_SPLIT_PARTS;
static_assert(SP_TILE0 == SP_THREADS, "one row per thread and plan tile");
static_assert(SP_MAX_PARTS <= 32, "one lane per destination");

// ---------------------------------------------------------------- plan
// One WARP per tile of 256 rows (8 rows per lane, all loads independent), no block barriers
// after the splitters are staged: counts come from warp ballots.
//
// A key is first compared on a PREFIX only -- the leading 4 bytes of a fixed_size_binary key
// (two 16-bit loads instead of five for the 10-byte GraySort key), the whole 64-bit word for
// the numeric kinds.  The prefix decides the destination unless it equals the prefix of the
// splitter next to it; only such keys (a handful per million for spread keys) load the rest.
template <int KIND>
__device__ __forceinline__ uint64_t load_key_prefix(const void* col, uint32_t width, uint64_t i, bool desc) {
  if (KIND != KIND_FSB) {
    uint64_t h;
    uint32_t l;
    load_key<KIND>(col, width, i, desc, h, l);
    return h;
  }
  const uint8_t* p = static_cast<const uint8_t*>(col) + i * width;
  uint32_t v = 0;
  if ((width & 1) == 0 && width >= 4) {
    const uint16_t* q = reinterpret_cast<const uint16_t*>(p);
    const uint32_t a = __ldg(q), b = __ldg(q + 1);  // little-endian pairs
    v = ((a & 0xFF) << 24) | ((a >> 8) << 16) | ((b & 0xFF) << 8) | (b >> 8);
  } else {
#pragma unroll
    for (int b = 0; b < 4; ++b) v = (v << 8) | (b < (int)width ? __ldg(p + b) : 0u);
  }
  if (desc) v = ~v & (width >= 4 ? 0xFFFFFFFFu : ~((1u << (32 - 8 * width)) - 1u));
  return static_cast<uint64_t>(v) << 32;
}

// One full tile of 10-byte keys by one warp; see the call site.  The key of row j of a lane sits at
// byte 10 j of the lane's 80 bytes: at a word boundary for even j, two bytes into a word for odd j.
template <int NG>  // groups of four destinations that can occur: nb + 1 <= 4 * NG
__device__ __forceinline__ void split_plan_tile_fsb10(const void* __restrict__ col, const uint8_t* __restrict__ valid,
                                                      uint32_t t, const uint64_t* s_bh, const uint32_t* s_bl,
                                                      uint32_t nb, bool desc, uint8_t* __restrict__ dest_out,
                                                      uint32_t* __restrict__ tile_counts, uint32_t ntiles,
                                                      uint32_t lane) {
  constexpr int ROWS = SP_TILE0 / 32;  // 8
  static_assert(ROWS == 8, "a lane's rows are 80 bytes and its destinations one 8-byte word");
  const size_t row0 = static_cast<size_t>(t) * SP_TILE0 + lane * ROWS;
  const uint4* q = reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(col) + row0 * 10);
  uint32_t w[20];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const uint4 v = __ldg(q + k);
    w[4 * k] = v.x;
    w[4 * k + 1] = v.y;
    w[4 * k + 2] = v.z;
    w[4 * k + 3] = v.w;
  }
  unsigned long long vmask = ~0ull;  // byte j != 0: row j is not null
  if (valid != nullptr) vmask = *reinterpret_cast<const unsigned long long*>(valid + row0);
  const uint32_t nparts = nb + 1;
  unsigned long long dest8 = 0;
  unsigned long long acc[NG];  // per-lane counts, four 16-bit fields per word
#pragma unroll
  for (int g = 0; g < NG; ++g) acc[g] = 0;
#pragma unroll
  for (int j = 0; j < ROWS; ++j) {
    const int k = (10 * j) / 4;
    uint32_t h1, h0, l;
    if (j % 2 == 0) {
      h1 = __byte_perm(w[k], 0, 0x0123);
      h0 = __byte_perm(w[k + 1], 0, 0x0123);
      l = __byte_perm(w[k + 2], 0, 0x4401);
    } else {
      h1 = __byte_perm(w[k], w[k + 1], 0x2345);
      h0 = __byte_perm(w[k + 1], w[k + 2], 0x2345);
      l = __byte_perm(w[k + 2], 0, 0x4423);
    }
    uint64_t kh = (static_cast<uint64_t>(h1) << 32) | h0;
    uint32_t kl = l;
    if (desc) {  // the encoding of load_key<KIND_FSB>
      kh = ~kh;
      kl = ~kl & 0xFFFFu;
    }
    // asc: number of splitters <= key; desc: number of splitters strictly before the key (in
    // the encoded order) -- the rule of the general path below
    uint32_t a = 0, b = nb;
    while (a < b) {
      const uint32_t mid = (a + b) >> 1;
      const uint64_t sh = s_bh[mid];
      bool before = sh < kh;
      if (sh == kh) before = desc ? s_bl[mid] < kl : s_bl[mid] <= kl;
      if (before) a = mid + 1;
      else b = mid;
    }
    const uint32_t d = ((vmask >> (8 * j)) & 0xFFu) ? a : nb;  // null keys travel to the last partition
    dest8 |= static_cast<unsigned long long>(d) << (8 * j);
    const uint32_t grp = d >> 2, sh = 16 * (d & 3);
#pragma unroll
    for (int g = 0; g < NG; ++g)  // arithmetic, not `if`: keeps acc[] in registers
      acc[g] += static_cast<unsigned long long>(grp == (uint32_t)g) << sh;
  }
  *reinterpret_cast<unsigned long long*>(dest_out + row0) = dest8;
  uint32_t mine = 0;
#pragma unroll
  for (int g = 0; g < NG; ++g) {
    // a field holds at most 256 after the reduction: 16 bits are enough
    const uint32_t lo = __reduce_add_sync(0xffffffffu, static_cast<uint32_t>(acc[g]));
    const uint32_t hi = __reduce_add_sync(0xffffffffu, static_cast<uint32_t>(acc[g] >> 32));
    if ((lane >> 2) == (uint32_t)g) {
      const uint32_t pair = (lane & 2) ? hi : lo;
      mine = (pair >> (16 * (lane & 1))) & 0xFFFFu;
    }
  }
  if (lane < nparts) tile_counts[static_cast<size_t>(lane) * (ntiles + 1) + t] = mine;
}

template <int KIND>
__global__ void __launch_bounds__(SP_THREADS)
    split_plan_kernel(const void* __restrict__ col, uint32_t width, const uint8_t* __restrict__ valid,
                      uint32_t n, const void* __restrict__ bounds, uint32_t nb, bool desc,
                      uint8_t* __restrict__ dest_out, uint32_t* __restrict__ tile_counts, uint32_t ntiles) {
  __shared__ uint64_t s_bh[SP_MAX_PARTS];
  __shared__ uint32_t s_bl[SP_MAX_PARTS];
  __shared__ uint64_t s_bp[SP_MAX_PARTS];  // the splitters' prefixes
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t nparts = nb + 1;
  constexpr uint64_t PMASK = KIND == KIND_FSB ? 0xFFFFFFFF00000000ull : ~0ull;
  if (tid < nb) {
    uint64_t h;
    uint32_t l;
    load_key<KIND>(bounds, width, tid, desc, h, l);
    s_bh[tid] = h;
    s_bl[tid] = l;
    s_bp[tid] = h & PMASK;
  }
  __syncthreads();
  constexpr int PER = SP_TILE0 / 32;
  const uint32_t warps = gridDim.x * (SP_THREADS / 32);
  // 10-byte keys (the GraySort key), full tiles, 16-byte aligned column: a lane takes 8
  // CONSECUTIVE rows -- 80 bytes, five 16-byte loads, the whole key in registers -- and writes
  // their 8 destination bytes with one store (split_plan_tile_fsb10 below).
  const bool fast10 = KIND == KIND_FSB && width == 10 && (reinterpret_cast<uintptr_t>(col) & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(dest_out) & 7) == 0 &&
                      (valid == nullptr || (reinterpret_cast<uintptr_t>(valid) & 7) == 0);
  for (uint32_t t = blockIdx.x * (SP_THREADS / 32) + warp; t < ntiles; t += warps) {
    if (fast10 && static_cast<uint64_t>(t + 1) * SP_TILE0 <= n) {
      if (nb < 8) split_plan_tile_fsb10<2>(col, valid, t, s_bh, s_bl, nb, desc, dest_out, tile_counts, ntiles, lane);
      else if (nb < 16) split_plan_tile_fsb10<4>(col, valid, t, s_bh, s_bl, nb, desc, dest_out, tile_counts, ntiles, lane);
      else split_plan_tile_fsb10<8>(col, valid, t, s_bh, s_bl, nb, desc, dest_out, tile_counts, ntiles, lane);
      continue;
    }
    const uint32_t base = t * SP_TILE0 + lane;
    uint64_t kp[PER];
    uint32_t d[PER];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const uint32_t i = base + j * 32;
      kp[j] = 0;
      d[j] = 0xffffffffu;
      if (i < n) {
        d[j] = 0xfffffffeu;  // "to be searched"
        if (valid != nullptr && valid[i] == 0) d[j] = nb;  // null keys travel to the last partition
        else kp[j] = load_key_prefix<KIND>(col, width, i, desc);
      }
    }
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const uint32_t i = base + j * 32;
      if (d[j] == 0xfffffffeu) {
        // lo = #{splitter prefixes < key prefix}, hi = #{splitter prefixes <= key prefix}
        uint32_t lo = 0, hi = nb;
        while (lo < hi) {
          const uint32_t mid = (lo + hi) >> 1;
          if (s_bp[mid] < kp[j]) lo = mid + 1;
          else hi = mid;
        }
        uint32_t up = lo;
        while (up < nb && s_bp[up] == kp[j]) ++up;
        if (up != lo) {
          // some splitters share the prefix: the full key decides among [lo, up]
          uint64_t kh;
          uint32_t kl;
          load_key<KIND>(col, width, i, desc, kh, kl);
          uint32_t b = lo;
          while (b < up) {
            // asc: count splitters <= key; desc: count splitters strictly before key (encoded)
            const bool go_right = desc ? key_less(s_bh[b], s_bl[b], kh, kl) : !key_less(kh, kl, s_bh[b], s_bl[b]);
            if (!go_right) break;
            ++b;
          }
          d[j] = b;
        } else {
          d[j] = lo;  // no splitter shares the prefix: every splitter below it is <= the key
        }
      }
      if (i < n) dest_out[i] = static_cast<uint8_t>(d[j]);
    }
    uint32_t mine = 0;
    for (uint32_t b = 0; b < nparts; ++b) {
      uint32_t c = 0;
#pragma unroll
      for (int j = 0; j < PER; ++j) c += __popc(__ballot_sync(0xffffffffu, d[j] == b));
      if (lane == b) mine = c;
    }
    if (lane < nparts) tile_counts[static_cast<size_t>(lane) * (ntiles + 1) + t] = mine;
  }
}

// Exclusive scan over tiles of every destination's counts, in place, as reduce-then-scan:
//   split_seg_sum_kernel   one CTA per (segment of SCAN_SEG tiles, destination): segment total
//   split_seg_scan_kernel  one CTA per destination: exclusive scan of its segment totals;
//                          entry ntiles of the table and counts_out get the grand total
//   split_seg_apply_kernel the same CTAs as the first kernel: exclusive scan inside the segment
//                          (8 consecutive tiles per thread) plus the segment's offset
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_PER = 8;
constexpr int SCAN_SEG = SCAN_THREADS * SCAN_PER;

__device__ __forceinline__ uint32_t scan_block_exclusive(uint32_t v, uint32_t* s_warp, uint32_t* total) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t before = 0, all = 0;
  for (uint32_t w = 0; w < blockDim.x / 32; ++w) {
    const uint32_t c = s_warp[w];
    if (w < warp) before += c;
    all += c;
  }
  if (total) *total = all;
  return before + incl - v;
}

__device__ __forceinline__ void scan_load(const uint32_t* row, uint32_t ntiles, uint32_t first, uint32_t (&c)[SCAN_PER]) {
#pragma unroll
  for (int k = 0; k < SCAN_PER; ++k) c[k] = first + k < ntiles ? row[first + k] : 0u;
}

__global__ void __launch_bounds__(SCAN_THREADS)
    split_seg_sum_kernel(const uint32_t* __restrict__ tile_counts, uint32_t ntiles, uint32_t nsegs,
                         uint32_t* __restrict__ seg_sums) {
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  const uint32_t* row = tile_counts + static_cast<size_t>(blockIdx.y) * (ntiles + 1);
  uint32_t c[SCAN_PER];
  scan_load(row, ntiles, blockIdx.x * SCAN_SEG + threadIdx.x * SCAN_PER, c);
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_PER; ++k) sum += c[k];
  uint32_t total;
  scan_block_exclusive(sum, s_warp, &total);
  if (threadIdx.x == 0) seg_sums[blockIdx.y * nsegs + blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024)
    split_seg_scan_kernel(uint32_t* __restrict__ seg_sums, uint32_t nsegs, uint32_t* __restrict__ tile_counts,
                          uint32_t ntiles, uint64_t* __restrict__ counts_out) {
  __shared__ uint32_t s_warp[32];
  uint32_t* row = seg_sums + static_cast<size_t>(blockIdx.x) * nsegs;
  const uint32_t per = (nsegs + 1023) / 1024;
  const uint32_t lo = min(threadIdx.x * per, nsegs), hi = min(lo + per, nsegs);
  uint32_t sum = 0;
  for (uint32_t i = lo; i < hi; ++i) sum += row[i];
  uint32_t total;
  uint32_t run = scan_block_exclusive(sum, s_warp, &total);
  for (uint32_t i = lo; i < hi; ++i) {
    const uint32_t c = row[i];
    row[i] = run;
    run += c;
  }
  if (threadIdx.x == 0) {
    tile_counts[static_cast<size_t>(blockIdx.x) * (ntiles + 1) + ntiles] = total;
    counts_out[blockIdx.x] = total;
  }
}

__global__ void __launch_bounds__(SCAN_THREADS)
    split_seg_apply_kernel(uint32_t* __restrict__ tile_counts, uint32_t ntiles, uint32_t nsegs,
                           const uint32_t* __restrict__ seg_sums) {
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  uint32_t* row = tile_counts + static_cast<size_t>(blockIdx.y) * (ntiles + 1);
  const uint32_t first = blockIdx.x * SCAN_SEG + threadIdx.x * SCAN_PER;
  uint32_t c[SCAN_PER];
  scan_load(row, ntiles, first, c);
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_PER; ++k) sum += c[k];
  uint32_t run = scan_block_exclusive(sum, s_warp, nullptr) + seg_sums[blockIdx.y * nsegs + blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_PER; ++k) {
    if (first + k < ntiles) row[first + k] = run;
    run += c[k];
  }
}

// ---------------------------------------------------------------- split
struct SplitBases {
  unsigned char* p[SP_MAX_PARTS];
};

// generic (any alignment, any width) row copy inside shared memory
__device__ __forceinline__ void copy_row_bytes(uint8_t* __restrict__ sm, uint32_t s, uint32_t d, uint32_t W) {
  for (uint32_t b = 0; b < W; ++b) sm[d + b] = sm[s + b];
}

// one aligned store of 1 << k bytes (k = 0..3), shared -> global
__device__ __forceinline__ void store_piece(uint8_t* g, const uint8_t* sm, uint32_t k) {
  if (k == 0) *g = *sm;
  else if (k == 1) *reinterpret_cast<uint16_t*>(g) = *reinterpret_cast<const uint16_t*>(sm);
  else if (k == 2) *reinterpret_cast<uint32_t*>(g) = *reinterpret_cast<const uint32_t*>(sm);
  else *reinterpret_cast<uint2*>(g) = *reinterpret_cast<const uint2*>(sm);
}

template <int K>  // tile = K * 256 rows
__global__ void __launch_bounds__(SP_THREADS)
    split_rows_kernel(const uint8_t* __restrict__ src, uint32_t W, uint32_t n, const uint8_t* __restrict__ dest,
                      const uint32_t* __restrict__ tile_offsets, uint32_t ntiles0, uint32_t nparts,
                      SplitBases bases, uint32_t tiles, bool even_path, uint32_t dest_mask) {
  constexpr uint32_t T = K * SP_TILE0;
  constexpr uint32_t CHUNKS = T / 32;
  extern __shared__ __align__(128) uint8_t sp_sm[];
  // layout: bars | chunk counts | A[2] (tiles as they lie in the input) | B[2] (spans)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sp_sm);                  // [2]
  uint32_t* s_cc = reinterpret_cast<uint32_t*>(sp_sm + 16);             // [CHUNKS][32]
  const uint32_t a_off0 = 16 + CHUNKS * 32 * 4 + 32 * 4;                // multiple of 16
  const uint32_t a_bytes = T * W;                                       // multiple of 16
  const uint32_t a_stride = a_bytes + 16;                               // + spare for the row copy's look-ahead word
  const uint32_t b_off0 = a_off0 + 2 * a_stride;
  const uint32_t b_stride = a_bytes + 32 * SP_MAX_PARTS;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const bool src_aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
  const size_t my_table = static_cast<size_t>(lane) * (ntiles0 + 1);    // lane b < nparts: its row of the table

  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    mbar_fence_init();
  }
  __syncthreads();

  const uint32_t first = blockIdx.x;
  const uint32_t my_tiles = tiles > first ? (tiles - first + gridDim.x - 1) / gridDim.x : 0;
  auto tile_of = [&](uint32_t k) { return first + k * gridDim.x; };

  // tile k of this CTA -> A[k & 1]; called by every thread (the bulk path is one thread's work)
  auto load_tile = [&](uint32_t k) {
    const uint32_t row0 = tile_of(k) * T;
    const uint32_t rows = min(T, n - row0);
    const uint8_t* g = src + static_cast<uint64_t>(row0) * W;
    const uint32_t a_off = a_off0 + (k & 1u) * a_stride;
    if (rows == T && src_aligned) {
      if (tid == 0) {
        mbar_arrive_expect_tx(&bars[k & 1u], a_bytes);
        bulk_g2s(sp_sm + a_off, g, a_bytes, &bars[k & 1u]);
      }
    } else {  // ragged last tile or an unaligned column: plain loads
      const uint32_t nbytes = rows * W;
      for (uint32_t b = tid; b < nbytes; b += SP_THREADS) sp_sm[a_off + b] = g[b];
      __syncthreads();
      if (tid == 0) mbar_arrive_expect_tx(&bars[k & 1u], 0);
    }
  };
  // what the ranking of tile k needs from global memory, fetched one tile ahead into registers:
  // the destinations of this thread's rows and (thread b < nparts) the table entries around the tile
  uint32_t nd[K], noff0 = 0, noff1 = 0;
  auto prefetch = [&](uint32_t k) {
    const uint32_t t = tile_of(k);
    const uint32_t row0 = t * T;
    const uint32_t rows = min(T, n - row0);
#pragma unroll
    for (int q = 0; q < K; ++q) {
      const uint32_t i = q * SP_TILE0 + tid;
      nd[q] = i < rows ? dest[row0 + i] : 0xffffffffu;
    }
    if (lane < nparts) {  // every warp lays the spans out for itself
      noff0 = tile_offsets[my_table + t * K];
      noff1 = tile_offsets[my_table + min(t * K + K, ntiles0)];
    }
  };

  if (my_tiles) {
    prefetch(0);
    load_tile(0);
    if (my_tiles > 1) load_tile(1);
  }
  for (uint32_t it = 0; it < my_tiles; ++it) {
    const uint32_t rows = min(T, n - tile_of(it) * T);
    const uint32_t a_off = a_off0 + (it & 1u) * a_stride;
    const uint32_t b_off = b_off0 + (it & 1u) * b_stride;
    uint32_t d[K], rk[K];
#pragma unroll
    for (int q = 0; q < K; ++q) d[q] = nd[q];
    const uint32_t off0 = noff0, off1 = noff1;
    if (it + 1 < my_tiles) prefetch(it + 1);  // lands while this tile is being processed

    // ---- stable ranks (row i of the tile = chunk i/32, lane i%32)
#pragma unroll
    for (int q = 0; q < K; ++q) {
      uint32_t mine = 0;
      rk[q] = 0;
      for (uint32_t b = 0; b < nparts; ++b) {
        if (!(dest_mask >> b & 1u)) continue;  // rows of destinations outside the mask stay behind
        const uint32_t m = __ballot_sync(0xffffffffu, d[q] == b);
        if (d[q] == b) rk[q] = __popc(m & lt_mask);
        if (lane == b) mine = __popc(m);
      }
      s_cc[(q * 8 + warp) * 32 + lane] = mine;
    }
    // the bulk stores that read B[it & 1] two tiles ago must have drained (lane 0 of every
    // warp ships the spans of its destinations, one bulk group per tile)
    if (lane == 0 && it >= 2) bulk_wait_read<1>();
    __syncthreads();
    // Every warp lays the spans out for itself (no second barrier, no serial section): lane b
    // sums destination b's chunk counts -- those of earlier chunks give the offset of this
    // warp's rows inside the span, all of them the span's length -- and a warp scan over the
    // padded lengths places the spans in B with the byte alignment of their global address.
    uint32_t before[K], total = 0;
#pragma unroll
    for (uint32_t c = 0; c < CHUNKS; ++c) {
      if ((c & 7u) == warp) before[c >> 3] = total;
      total += s_cc[c * 32 + lane];
    }
    uint32_t span_len = 0, res = 0, padded = 0;
    uint8_t* span_g = nullptr;
    if (lane < nparts && total) {  // total == off1 - off0 by construction of the table
      span_g = bases.p[lane] + static_cast<uint64_t>(off0) * W;
      span_len = (off1 - off0) * W;
      res = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(span_g) & 15);
      padded = (res + span_len + 15u) & ~15u;
    }
    uint32_t incl = padded;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)lane >= o) incl += v;
    }
    const uint32_t span_start = incl - padded + res;  // lane b: byte offset of span b's rows in B
    uint32_t boff[K];
#pragma unroll
    for (int q = 0; q < K; ++q) {
      const uint32_t base = __shfl_sync(0xffffffffu, span_start + before[q] * W, d[q] & 31u);
      boff[q] = base + rk[q] * W;
    }

    // ---- the tile has arrived: move every row to its place in its span
    mbar_wait(&bars[it & 1u], (it >> 1) & 1u);
#pragma unroll
    for (int q = 0; q < K; ++q) {
      const uint32_t i = q * SP_TILE0 + tid;
      if (i < rows && d[q] < nparts && (dest_mask >> d[q] & 1u)) {
        if (even_path) copy_row_even(sp_sm, a_off + i * W, b_off + boff[q], W);
        else copy_row_bytes(sp_sm, a_off + i * W, b_off + boff[q], W);
      }
    }
    fence_proxy_async_smem();
    __syncthreads();  // spans complete; A[it & 1] is free for the tile after next
    if (it + 2 < my_tiles) load_tile(it + 2);

    // ---- ship the spans: warp w takes destinations w, w + 8, ...; its lanes copy the few
    // bytes before and after the 16-byte-aligned body, lane 0 issues the body's bulk store
    for (uint32_t b = warp; b < nparts; b += SP_THREADS / 32) {
      const uint32_t start = __shfl_sync(0xffffffffu, span_start, b);
      const uint32_t len = __shfl_sync(0xffffffffu, span_len, b);
      const unsigned long long ga = __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(span_g), b);
      if (len == 0) continue;
      uint8_t* g = reinterpret_cast<uint8_t*>(ga);
      const uint8_t* sdat = sp_sm + b_off + start;
      const uint32_t gap = (16u - (start & 15u)) & 15u;
      const uint32_t head = min(len, gap);
      const uint32_t body = (len - head) & ~15u;
      const uint32_t tail = len - head - body;  // < 16
      if (head < gap) {  // a span shorter than the distance to the next 16-byte boundary: bytes
        if (lane < head) g[lane] = sdat[lane];
        continue;
      }
      // The bytes before and after the aligned body go out as the few widest aligned stores
      // that cover them (8/4/2/1 bytes: every remote store is a packet of its own on NVLink).
      // head < 16 ends on a 16-byte boundary: lane k (0..3) stores the piece of 1 << k bytes if
      // bit k of head is set, pieces in ascending size; the tail mirrors it from its start.
      if (lane < 4 && (head >> lane & 1u)) {
        const uint32_t o = head & ((1u << lane) - 1u);  // bytes of smaller pieces come first
        store_piece(g + o, sdat + o, lane);
      } else if (lane >= 4 && lane < 8 && (tail >> (lane - 4) & 1u)) {
        const uint32_t k = lane - 4;
        const uint32_t o = head + body + (tail & ~((2u << k) - 1u));  // larger pieces come first
        store_piece(g + o, sdat + o, k);
      }
      if (lane == 0 && body) bulk_s2g(g + head, sdat + head, body);
    }
    if (lane == 0) bulk_commit();
  }
  if (lane == 0) bulk_wait<0>();
}

static uint32_t g_split_k = 0;     // ds_debug_set_split_tile: force tile = k * 256 rows (0 = by row width)
static uint32_t g_split_ctas = 0;  // ds_debug_set_split_ctas: persistent CTAs per SM (0 = default: up to 3)

template <int K>
static cudaError_t launch_split(const void* src, uint32_t W, uint32_t n, const uint8_t* dest,
                                const uint32_t* tile_offsets, uint32_t nparts, const SplitBases& bases,
                                bool even_path, uint32_t dest_mask, uint32_t max_ctas, cudaStream_t s) {
  constexpr uint32_t T = K * SP_TILE0;
  const uint32_t ntiles0 = (n + SP_TILE0 - 1) / SP_TILE0;
  const uint32_t tiles = (n + T - 1) / T;
  const uint32_t smem = 16 + (T / 32) * 32 * 4 + 32 * 4 + 2 * (T * W + 16) + 2 * (T * W + 32 * SP_MAX_PARTS) + 16;
  auto k = split_rows_kernel<K>;
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (e != cudaSuccess) return e;
  // as many persistent CTAs per SM as fit (up to 3) unless the caller bounds the grid
  uint32_t per_sm = g_split_ctas ? g_split_ctas : 3;
  const uint32_t fit = (224u * 1024) / (smem + 1024);
  if (per_sm > fit) per_sm = fit ? fit : 1;
  uint32_t blocks = static_cast<uint32_t>(sm_count()) * per_sm;
  if (max_ctas && blocks > max_ctas) blocks = max_ctas;
  if (blocks > tiles) blocks = tiles;
  note_launch();
  k<<<blocks, SP_THREADS, smem, s>>>(static_cast<const uint8_t*>(src), W, n, dest, tile_offsets, ntiles0, nparts,
                                     bases, tiles, even_path, dest_mask);
  return cudaGetLastError();
}

}  // namespace ds

extern "C" {

ds_status ds_split_plan(int32_t key_kind, const void* col, uint32_t width, const uint8_t* valid, uint64_t n,
                        const void* boundaries, uint32_t nb, int32_t descending, uint8_t* dest_out,
                        uint32_t* tile_offsets_out, uint64_t* counts_out, void* stream) {
  using namespace ds;
  if (n > DS_MAX_ROWS) {
    set_error("%s: n=%llu exceeds DS_MAX_ROWS", __func__, (unsigned long long)n);
    return DS_ERR_TOO_LARGE;
  }
  if (nb + 1 > DS_MAX_SPLIT_PARTS) {
    set_error("%s: %u destinations exceed DS_MAX_SPLIT_PARTS (%d)", __func__, nb + 1, DS_MAX_SPLIT_PARTS);
    return DS_ERR_UNSUPPORTED;
  }
  if (key_kind == DS_KEY_FSB ? (width < 1 || width > DS_MAX_KEY_BYTES) : false) {
    set_error("%s: key width %u unsupported", __func__, width);
    return DS_ERR_UNSUPPORTED;
  }
  if (!tile_offsets_out || !counts_out || (n > 0 && (!col || !dest_out)) || (nb > 0 && !boundaries)) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint32_t n32 = static_cast<uint32_t>(n);
  const uint32_t ntiles = (n32 + SP_TILE0 - 1) / SP_TILE0;
  if (n32) {
    uint32_t blocks = static_cast<uint32_t>(sm_count()) * 8;
    if (blocks > (ntiles + 7) / 8) blocks = (ntiles + 7) / 8;
    note_launch();
    const bool desc = descending != 0;
    switch (key_kind) {
      case DS_KEY_I64:
        split_plan_kernel<KIND_I64><<<blocks, SP_THREADS, 0, s>>>(col, 8, valid, n32, boundaries, nb, desc, dest_out, tile_offsets_out, ntiles);
        break;
      case DS_KEY_U64:
        split_plan_kernel<KIND_U64><<<blocks, SP_THREADS, 0, s>>>(col, 8, valid, n32, boundaries, nb, desc, dest_out, tile_offsets_out, ntiles);
        break;
      case DS_KEY_F64:
        split_plan_kernel<KIND_F64><<<blocks, SP_THREADS, 0, s>>>(col, 8, valid, n32, boundaries, nb, desc, dest_out, tile_offsets_out, ntiles);
        break;
      case DS_KEY_FSB:
        split_plan_kernel<KIND_FSB><<<blocks, SP_THREADS, 0, s>>>(col, width, valid, n32, boundaries, nb, desc, dest_out, tile_offsets_out, ntiles);
        break;
      default:
        set_error("%s: unknown key kind %d", __func__, key_kind);
        return DS_ERR_INVALID_ARG;
    }
    DS_LAUNCH_CHECK();
  }
  // the segment totals of the scan live in the tail of the table's allocation
  // (ds_split_table_words accounts for them)
  const uint32_t nsegs = (ntiles + SCAN_SEG - 1) / SCAN_SEG;
  uint32_t* seg_sums = tile_offsets_out + static_cast<size_t>(nb + 1) * (ntiles + 1);
  if (nsegs) {
    note_launch();
    split_seg_sum_kernel<<<dim3(nsegs, nb + 1), SCAN_THREADS, 0, s>>>(tile_offsets_out, ntiles, nsegs, seg_sums);
    DS_LAUNCH_CHECK();
  }
  note_launch();
  split_seg_scan_kernel<<<nb + 1, 1024, 0, s>>>(seg_sums, nsegs, tile_offsets_out, ntiles, counts_out);
  DS_LAUNCH_CHECK();
  if (nsegs) {
    note_launch();
    split_seg_apply_kernel<<<dim3(nsegs, nb + 1), SCAN_THREADS, 0, s>>>(tile_offsets_out, ntiles, nsegs, seg_sums);
    DS_LAUNCH_CHECK();
  }
  return DS_OK;
}

uint64_t ds_split_table_words(uint64_t n, uint32_t nparts) {
  const uint64_t ntiles = (n + ds::SP_TILE0 - 1) / ds::SP_TILE0;
  const uint64_t nsegs = (ntiles + ds::SCAN_SEG - 1) / ds::SCAN_SEG;
  return static_cast<uint64_t>(nparts) * (ntiles + 1 + nsegs);
}

ds_status ds_split_rows(const void* src, uint32_t row_bytes, uint64_t n, const uint8_t* dest,
                        const uint32_t* tile_offsets, uint32_t nparts, void* const* dst_bases,
                        uint32_t dest_mask, uint32_t max_ctas, void* stream) {
  using namespace ds;
  if (n == 0) return DS_OK;
  if (n > DS_MAX_ROWS) {
    set_error("%s: n=%llu exceeds DS_MAX_ROWS", __func__, (unsigned long long)n);
    return DS_ERR_TOO_LARGE;
  }
  if (nparts == 0 || nparts > DS_MAX_SPLIT_PARTS) {
    set_error("%s: %u destinations unsupported (1..%d)", __func__, nparts, DS_MAX_SPLIT_PARTS);
    return DS_ERR_UNSUPPORTED;
  }
  if (row_bytes == 0 || row_bytes > DS_SPLIT_MAX_ROW_BYTES) {
    set_error("%s: row_bytes=%u unsupported (1..%d)", __func__, row_bytes, DS_SPLIT_MAX_ROW_BYTES);
    return DS_ERR_UNSUPPORTED;
  }
  if (!src || !dest || !tile_offsets || !dst_bases) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  if (dest_mask == 0) dest_mask = 0xffffffffu;
  SplitBases bases{};
  uintptr_t bits = reinterpret_cast<uintptr_t>(src) | row_bytes;
  for (uint32_t d = 0; d < nparts; ++d) {
    bases.p[d] = static_cast<unsigned char*>(dst_bases[d]);
    if (dest_mask >> d & 1u) {
      if (dst_bases[d] == nullptr) {
        set_error("%s: dst_bases[%u] is NULL", __func__, d);
        return DS_ERR_INVALID_ARG;
      }
      bits |= reinterpret_cast<uintptr_t>(dst_bases[d]);
    }
  }
  const bool even_path = (bits & 1) == 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint32_t n32 = static_cast<uint32_t>(n);
  const uint8_t* dst8 = dest;
  cudaError_t e;
  // tile of K * 256 rows: the largest whose four shared-memory images stay under ~100 KB
  const uint32_t per256 = 4u * SP_TILE0 * row_bytes;
  uint32_t kk = g_split_k;
  if (kk && per256 * kk > 220 * 1024) kk = 0;
  if (kk == 2) e = launch_split<2>(src, row_bytes, n32, dst8, tile_offsets, nparts, bases, even_path, dest_mask, max_ctas, s);
  else if (kk == 4) e = launch_split<4>(src, row_bytes, n32, dst8, tile_offsets, nparts, bases, even_path, dest_mask, max_ctas, s);
  else if (kk == 1) e = launch_split<1>(src, row_bytes, n32, dst8, tile_offsets, nparts, bases, even_path, dest_mask, max_ctas, s);
  else if (per256 * 8 <= 100 * 1024) e = launch_split<8>(src, row_bytes, n32, dst8, tile_offsets, nparts, bases, even_path, dest_mask, max_ctas, s);
  else if (per256 * 4 <= 100 * 1024) e = launch_split<4>(src, row_bytes, n32, dst8, tile_offsets, nparts, bases, even_path, dest_mask, max_ctas, s);
  else if (per256 * 2 <= 100 * 1024) e = launch_split<2>(src, row_bytes, n32, dst8, tile_offsets, nparts, bases, even_path, dest_mask, max_ctas, s);
  else e = launch_split<1>(src, row_bytes, n32, dst8, tile_offsets, nparts, bases, even_path, dest_mask, max_ctas, s);
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", __func__, cudaGetErrorString(e));
    return DS_ERR_CUDA;
  }
  return DS_OK;
}

// Tuning hook: tile of ds_split_rows = k * 256 rows (1, 2, 4; 0 = chosen from the row width).
void ds_debug_set_split_tile(uint32_t k) { ds::g_split_k = k; }

// Tuning hook: persistent CTAs per SM of ds_split_rows (0 = default).
void ds_debug_set_split_ctas(uint32_t per_sm) { ds::g_split_ctas = per_sm; }

}  // extern "C"
