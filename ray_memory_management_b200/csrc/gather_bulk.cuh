// Row gather for even-sized rows that are not a multiple of 16 bytes (the 90-byte GraySort
// payload), built on asynchronous copies instead of per-thread global loads and stores.
//
// A persistent CTA walks tiles of R output rows through a ring of STAGES shared-memory
// stages.  For every source row it fetches the 16-byte-aligned span that covers the row
// (W + misalignment rounded up to 16 bytes) into the row's slot of a stage:
//   FETCH_BULK    one cp.async.bulk per row (SASS UBLKCP), completion counted on the stage's
//                 mbarrier through expect_tx / complete_tx;
//   FETCH_LDGSTS  16-byte cp.async chunks with a 64-byte L2 fetch hint (SASS LDGSTS),
//                 consecutive lanes taking consecutive chunks of a row, completion by
//                 cp.async groups.
// While later stages are in flight, one thread per row copies the row from its slot, at
// whatever 2-byte alignment it arrived with, into the tile's packed output image in shared
// memory (aligned 32-bit words, byte_perm for the half-word shift), and the image leaves as
// ONE bulk store (cp.async.bulk shared -> global).  Global memory therefore only sees aligned
// 16-byte-multiple transfers issued by the SM's copy engine; the odd row alignment is
// confined to shared memory, and a row costs ~70 instructions instead of the ~270 of the
// per-granule kernels.  Rows whose aligned span would leave [src, src + src_bytes) (first /
// last rows of the buffer) are fetched byte by byte.
//
// Algorithmic bytes per row: 4 (index) + 2 * W.
#pragma once
#include "ds_async.cuh"
#include "ds_common.cuh"

namespace ds {

constexpr int GB_THREADS = 256;
constexpr int GB_MAX_STAGES = 4;
enum { FETCH_BULK = 0, FETCH_LDGSTS = 1 };

struct GatherBulkPlan {
  uint32_t rows;    // R: output rows per tile = threads per CTA (multiple of 32, <= GB_THREADS)
  uint32_t stages;  // ring depth
  uint32_t slot;    // bytes of shared memory per fetched row (odd multiple of 16)
  uint32_t chunks;  // 16-byte chunks per slot
  uint32_t smem;    // dynamic shared memory per CTA
  uint32_t ctas_per_sm;
};

__device__ __forceinline__ uint32_t gb_warp_sum(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Copies one row of W bytes (W even) between two locations of the CTA's dynamic shared memory
// given as byte offsets from its 16-byte-aligned base, both 2-byte aligned: aligned 32-bit
// stores, each word assembled from the two aligned source words that cover it.  May read (not
// use) one aligned word past the row's last byte.
__device__ __forceinline__ void copy_row_even(uint8_t* __restrict__ sm, uint32_t s, uint32_t d, uint32_t W) {
  uint32_t left = W;
  if (d & 2) {  // leading half word up to the next aligned word of the destination
    *reinterpret_cast<uint16_t*>(sm + d) = *reinterpret_cast<const uint16_t*>(sm + s);
    s += 2;
    d += 2;
    left -= 2;
  }
  const uint32_t sel = (s & 2) ? 0x5432u : 0x3210u;
  const uint32_t* w = reinterpret_cast<const uint32_t*>(sm) + (s >> 2);
  uint32_t* o = reinterpret_cast<uint32_t*>(sm) + (d >> 2);
  const uint32_t nw = left >> 2;
  uint32_t prev = w[0];
#pragma unroll 8
  for (uint32_t k = 0; k < nw; ++k) {
    const uint32_t cur = w[k + 1];
    o[k] = __byte_perm(prev, cur, sel);
    prev = cur;
  }
  if (left & 2)  // sel 0x5432: bytes 2..3 of the last word read; 0x3210: its bytes 0..1
    *reinterpret_cast<uint16_t*>(sm + d + (nw << 2)) = static_cast<uint16_t>(__byte_perm(prev, 0, sel));
}

// The source column may lie in several buffers (the blocks of a dataset): idx counts rows over
// their concatenation, source b holds rows [start[b], start[b+1]).
constexpr int GB_MAX_SOURCES = 64;
struct GatherSources {
  uint32_t nsrc;
  uint32_t start[GB_MAX_SOURCES + 1];
  const uint8_t* base[GB_MAX_SOURCES];
};
constexpr uint64_t GB_GUARDED = 1ull << 63;  // flag in a resolved address: fetch this row byte by byte

// One thread per row of a tile: blockDim.x == R.  MULTI = false: one source (srcs.base[0], rows
// [0, srcs.start[1])) -- its address and bounds stay in registers, no source table in shared memory.
template <int FETCH, bool MULTI>
__global__ void __launch_bounds__(GB_THREADS)
    gather_bulk_kernel(const GatherSources srcs, uint32_t W, const uint32_t* __restrict__ idx, uint32_t n,
                       uint8_t* __restrict__ dst, uint32_t R, uint32_t stages, uint32_t slot, uint32_t chunks,
                       uint32_t magic_c, uint32_t tiles) {
  extern __shared__ __align__(128) uint8_t gb_sm[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(gb_sm);                       // [GB_MAX_STAGES]
  uint64_t* s_base = reinterpret_cast<uint64_t*>(gb_sm + 64);                // [GB_MAX_SOURCES]
  uint32_t* s_start = reinterpret_cast<uint32_t*>(gb_sm + 64 + GB_MAX_SOURCES * 8);  // [GB_MAX_SOURCES + 1]
  constexpr uint32_t HDR = 64 + GB_MAX_SOURCES * 8 + (GB_MAX_SOURCES + 1) * 4 + 60;  // multiple of 64
  uint64_t* s_addr = reinterpret_cast<uint64_t*>(gb_sm + HDR);               // [R]: resolved rows of the tile fetched next
  uint32_t* rowoff = reinterpret_cast<uint32_t*>(gb_sm + HDR + R * 8);       // [stages][R]: byte offset of the row in its stage
  uint8_t* data = gb_sm + HDR + R * 8 + stages * R * 4;                      // [stages][R * slot] (+16 spare), 16-byte aligned
  const uint32_t image_bytes = R * W;                                        // multiple of 16 (R is)
  uint8_t* image = data + stages * R * slot + 16;                            // packed output tile
  (void)image_bytes;
  const uint32_t tid = threadIdx.x;
  const uint32_t lane = tid & 31;
  const uint32_t nsrc = srcs.nsrc;
  if (MULTI) {
    for (uint32_t b = tid; b < nsrc; b += blockDim.x) {
      s_base[b] = reinterpret_cast<uint64_t>(srcs.base[b]);
      s_start[b] = srcs.start[b];
    }
    if (tid == 0) s_start[nsrc] = srcs.start[nsrc];
    __syncthreads();
  }
  const uint64_t base0 = reinterpret_cast<uint64_t>(srcs.base[0]);
  const uint64_t safe_lo0 = (base0 + 15) & ~uint64_t(15);
  const uint64_t safe_hi0 = (base0 + static_cast<uint64_t>(srcs.start[1]) * W) & ~uint64_t(15);
  // row number over the concatenated sources -> address of the row; GB_GUARDED is set when the
  // 16-byte-aligned span around the row would leave its source buffer
  auto resolve = [&](uint32_t row) -> uint64_t {
    if (!MULTI) {
      const uint64_t a = base0 + static_cast<uint64_t>(row) * W;
      const bool safe = (a & ~uint64_t(15)) >= safe_lo0 && ((a + W + 15) & ~uint64_t(15)) <= safe_hi0;
      return safe ? a : (a | GB_GUARDED);
    }
    uint32_t b = 0;
    if (nsrc > 1) {
      uint32_t lo = 0, hi = nsrc;  // last b with start[b] <= row
      while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (s_start[mid] <= row) lo = mid;
        else hi = mid;
      }
      b = lo;
    }
    const uint64_t base = s_base[b];
    const uint64_t a = base + static_cast<uint64_t>(row - s_start[b]) * W;
    const uint64_t end = base + static_cast<uint64_t>(s_start[b + 1] - s_start[b]) * W;
    const uint64_t a0 = a & ~uint64_t(15), a1 = (a + W + 15) & ~uint64_t(15);
    const bool safe = a0 >= ((base + 15) & ~uint64_t(15)) && a1 <= (end & ~uint64_t(15));
    return safe ? a : (a | GB_GUARDED);
  };

  if (FETCH == FETCH_BULK) {
    if (tid == 0) {
      for (uint32_t s = 0; s < stages; ++s) mbar_init(&bars[s], R / 32);
      mbar_fence_init();
    }
    __syncthreads();
  }

  // `my` = resolved source row of output row t * R + tid, loaded one tile ahead of its use
  auto issue = [&](uint32_t t, uint32_t s, uint64_t my) {
    const uint32_t row0 = t * R;
    const uint32_t rows = min(R, n - row0);
    uint8_t* d = data + s * R * slot;
    uint32_t* ro = rowoff + s * R;
    if (FETCH == FETCH_BULK) {
      uintptr_t a0 = 0;
      uint32_t bytes = 0;
      if (tid < rows) {
        const uintptr_t a = static_cast<uintptr_t>(my & ~GB_GUARDED);
        const uint32_t sh = static_cast<uint32_t>(a & 15);
        a0 = a - sh;
        bytes = (sh + W + 15u) & ~15u;
        ro[tid] = tid * slot + sh;
        if (my & GB_GUARDED) {
          uint8_t* q = d + tid * slot + sh;
          for (uint32_t b = 0; b < W; ++b) q[b] = *reinterpret_cast<const uint8_t*>(a + b);
          bytes = 0;
        }
      }
      const uint32_t warp_bytes = gb_warp_sum(bytes);
      __syncwarp();  // the lanes' shared-memory stores are ordered before lane 0's release
      if (lane == 0) mbar_arrive_expect_tx(&bars[s], warp_bytes);
      __syncwarp();
      if (bytes) bulk_g2s(d + tid * slot, reinterpret_cast<const void*>(a0), bytes, &bars[s]);
    } else {
      // consecutive lanes take consecutive 16-byte chunks of a row: the source rows come from
      // shared memory (written by their owners before the last barrier)
      const uint32_t items = rows * chunks;
      for (uint32_t e = tid; e < items; e += R) {
        const uint32_t r = __umulhi(e, magic_c);
        const uint32_t c = e - r * chunks;
        const uint64_t ra = s_addr[r];
        const uintptr_t a = static_cast<uintptr_t>(ra & ~GB_GUARDED);
        const uint32_t sh = static_cast<uint32_t>(a & 15);
        if (c == 0) ro[r] = r * slot + sh;
        if (c * 16 >= sh + W) continue;  // chunk lies wholly behind the row
        const uintptr_t ac = a - sh + c * 16;
        if (!(ra & GB_GUARDED)) {
          ldgsts16_64B(d + r * slot + c * 16, reinterpret_cast<const void*>(ac));
        } else {  // first / last partial vector of the source buffer
          for (uint32_t b = 0; b < 16; ++b) {
            const uintptr_t p = ac + b;
            if (p >= a && p < a + W) d[r * slot + c * 16 + b] = *reinterpret_cast<const uint8_t*>(p);
          }
        }
      }
      ldgsts_commit();
    }
  };
  auto load_idx = [&](uint32_t k) -> uint64_t {  // resolved source row of this thread's row in my k-th tile
    const uint32_t row = (blockIdx.x + k * gridDim.x) * R + tid;
    return row < n ? resolve(ld_stream(idx + row)) : 0ull;
  };

  const uint32_t first = blockIdx.x;
  const uint32_t my_tiles = tiles > first ? (tiles - first + gridDim.x - 1) / gridDim.x : 0;
  // prologue: tiles 0 .. stages-2 are issued; `ahead` holds the source row for tile stages-1
  for (uint32_t k = 0; k + 1 < stages; ++k) {
    if (k < my_tiles) {
      const uint64_t my = load_idx(k);
      if (FETCH == FETCH_LDGSTS) {
        __syncthreads();
        s_addr[tid] = my;
        __syncthreads();
      }
      issue(first + k * gridDim.x, k, my);
    } else if (FETCH == FETCH_LDGSTS) {
      ldgsts_commit();  // keep the group count uniform
    }
  }
  uint64_t ahead = stages - 1 < my_tiles ? load_idx(stages - 1) : 0ull;
  if (FETCH == FETCH_LDGSTS) {
    __syncthreads();
    s_addr[tid] = ahead;
    __syncthreads();
  }
  for (uint32_t it = 0; it < my_tiles; ++it) {
    const uint32_t nxt = it + stages - 1;
    if (nxt < my_tiles) issue(first + nxt * gridDim.x, nxt % stages, ahead);
    else if (FETCH == FETCH_LDGSTS) ldgsts_commit();
    ahead = nxt + 1 < my_tiles ? load_idx(nxt + 1) : 0ull;  // in flight during this iteration
    const uint32_t s = it % stages;
    if (FETCH == FETCH_BULK) {
      mbar_wait(&bars[s], (it / stages) & 1u);
    } else {
      switch (stages) {  // all but the newest stages-1 groups are complete
        case 1: ldgsts_wait<0>(); break;
        case 2: ldgsts_wait<1>(); break;
        case 3: ldgsts_wait<2>(); break;
        default: ldgsts_wait<3>(); break;
      }
    }
    if (tid == 0 && it >= 1) bulk_wait_read<0>();  // the previous tile's image has been read out
    __syncthreads();

    // ---- realign: one thread per row, slot -> packed image
    const uint32_t t = first + it * gridDim.x;
    const uint32_t rows = min(R, n - t * R);
    if (tid < rows)
      copy_row_even(gb_sm, static_cast<uint32_t>(data - gb_sm) + s * R * slot + rowoff[s * R + tid],
                    static_cast<uint32_t>(image - gb_sm) + tid * W, W);
    if (FETCH == FETCH_LDGSTS) s_addr[tid] = ahead;  // read by the next iteration's issue
    fence_proxy_async_smem();
    __syncthreads();  // image complete; the stage may be refilled
    if (tid == 0) {
      uint8_t* out = dst + static_cast<uint64_t>(t) * R * W;
      const uint32_t nbytes = rows * W;
      const uint32_t body = nbytes & ~15u;
      if (body) bulk_s2g(out, image, body);
      bulk_commit();
      for (uint32_t b = body; b < nbytes; ++b) out[b] = image[b];  // ragged last tile only
    }
  }
  if (tid == 0) bulk_wait<0>();
}

// Tunables (ds_debug_set_gather_bulk): rows per tile, ring depth, CTAs per SM; 0 = default.
static uint32_t g_gb_rows = 0, g_gb_stages = 0, g_gb_ctas = 0;

static bool gather_bulk_plan(uint32_t W, GatherBulkPlan* p) {
  p->chunks = (W + 14 + 15) / 16;  // widest aligned span of a W-byte row that starts on an even address
  if ((p->chunks & 1) == 0) ++p->chunks;  // odd multiple of 16: neighbouring slots start in different banks
  p->slot = p->chunks * 16;
  p->stages = g_gb_stages ? g_gb_stages : 2;
  if (p->stages > GB_MAX_STAGES) p->stages = GB_MAX_STAGES;
  uint32_t rows = g_gb_rows ? g_gb_rows : 64;
  if (rows > GB_THREADS) rows = GB_THREADS;
  rows &= ~31u;
  const uint32_t budget = 110 * 1024;  // per CTA: stages + the output image
  // header (barriers + source table) + resolved rows + per-stage row offsets and slots + image
  auto need = [&](uint32_t r) { return 896u + r * 8 + p->stages * r * (p->slot + 4) + 16 + r * W; };
  while (rows > 32 && need(rows) > budget) rows -= 32;
  if (rows < 32 || need(rows) > 200 * 1024) return false;
  p->rows = rows;
  p->smem = need(rows);
  uint32_t ctas = (224u * 1024) / (p->smem + 1024);
  if (ctas > 8) ctas = 8;
  if (ctas < 1) ctas = 1;
  if (g_gb_ctas && g_gb_ctas < ctas) ctas = g_gb_ctas;
  p->ctas_per_sm = ctas;
  return true;
}

static bool gather_bulk_eligible(const void* src, uint32_t W, uint64_t src_rows, const void* dst) {
  return (W & 1) == 0 && (W & 15) != 0 && W >= 18 && W <= 1024 && src_rows > 0 &&
         (reinterpret_cast<uintptr_t>(src) & 1) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
}

static cudaError_t launch_gather_bulk(int fetch, const GatherSources& srcs, uint32_t W, const uint32_t* idx,
                                      uint32_t n, void* dst, cudaStream_t s, bool multi = true) {
  GatherBulkPlan p;
  if (!gather_bulk_plan(W, &p)) return cudaErrorInvalidConfiguration;
  const uint32_t tiles = (n + p.rows - 1) / p.rows;
  uint32_t blocks = static_cast<uint32_t>(sm_count()) * p.ctas_per_sm;
  if (blocks > tiles) blocks = tiles;
  const uint32_t magic_c = static_cast<uint32_t>((0x100000000ull + p.chunks - 1) / p.chunks);
  auto k = fetch == FETCH_BULK ? (multi ? gather_bulk_kernel<FETCH_BULK, true> : gather_bulk_kernel<FETCH_BULK, false>)
                               : (multi ? gather_bulk_kernel<FETCH_LDGSTS, true> : gather_bulk_kernel<FETCH_LDGSTS, false>);
  cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(p.smem));
  if (e != cudaSuccess) return e;
  ds::note_launch();
  k<<<blocks, p.rows, p.smem, s>>>(srcs, W, idx, n, static_cast<uint8_t*>(dst), p.rows, p.stages, p.slot,
                                   p.chunks, magic_c, tiles);
  return cudaGetLastError();
}

static cudaError_t launch_gather_bulk(int fetch, const void* src, uint32_t W, uint64_t src_rows,
                                      const uint32_t* idx, uint32_t n, void* dst, cudaStream_t s) {
  GatherSources srcs{};
  srcs.nsrc = 1;
  srcs.start[0] = 0;
  srcs.start[1] = static_cast<uint32_t>(src_rows);
  srcs.base[0] = static_cast<const uint8_t*>(src);
  return launch_gather_bulk(fetch, srcs, W, idx, n, dst, s, false);
}

}  // namespace ds
