// LSD "onesweep" radix sort of (key, uint32 row index) pairs for sm_100a.
//
// Replaces pyarrow.compute.sort_indices at
// python/ray/data/_internal/arrow_ops/transform_pyarrow.py:15,74 (reference).
//
// Data layout in HBM (all SoA, so every pass streams contiguous arrays):
//   hi[n]  uint64  radix-order encoding of the first <=8 key bytes
//   lo[n]  uint16  key bytes 8..9 of a fixed_size_binary(9|10) key (else absent)
//   idx[n] uint32  row index payload (generated, not loaded, by the first pass)
// Each array is double-buffered inside the caller's workspace.
//
// Kernels
//   encode_hist_kernel   reads the Arrow key column once, writes hi/lo and the 256-bin
//                        histograms of ALL digit positions (shared-memory atomics,
//                        one global flush per block).
//   scan_hist_kernel     exclusive scan per digit position -> global bin bases, and a
//                        "constant digit" flag per position (such passes are skipped).
//   onesweep_pass_kernel_v3 (onesweep_v3.cuh) one 8-bit pass: warp-striped coalesced loads,
//                        per-warp digit counts published to the look-back array before
//                        ranking, peers of a digit found through a per-warp shared-memory
//                        table (stable), per-digit decoupled look-back across tiles (tile ids
//                        handed out by an atomic counter, so a tile only ever waits on tiles
//                        that are already resident), shared-memory exchange, coalesced scatter.
//                        Algorithmic bytes per key per pass: 2*(8[+2]+4) = 24 / 28.
//   bucket_local_sort_kernel (bucket_sort.cuh) finishes a sort whose two leading bytes have
//                        been sorted globally: one CTA per 16-bit bucket in shared memory.
//   decode_kernel        sorted hi/lo -> sorted key column in its Arrow form.
#include <atomic>
#include <utility>
#include <vector>

#include "ds_common.cuh"
#include "ds_keys.cuh"

namespace ds {

constexpr int RADIX = 256;
constexpr uint32_t LB_AGGREGATE = 1u << 30;
constexpr uint32_t LB_INCLUSIVE = 2u << 30;
constexpr uint32_t LB_VALUE = (1u << 30) - 1;
constexpr int MAX_PASSES = 10;

struct PassList {
  int count;
  int shift[MAX_PASSES];
  int from_lo[MAX_PASSES];
};

// Keys that occupy a SUB-RANGE of the key space (one rank's share of a range-partitioned
// dataset, clustered ids, time stamps): the hybrid plan wants 65536 evenly filled buckets, which
// the two leading bytes of such keys do not give.  The encode kernel then stores
//   hi' = (hi - base) << s        (monotone and injective on [base, base + span))
// instead of hi, and every kernel of the plan takes its bucket / bin digits from the 28-bit
//   G = umul64hi(hi', R),  R = floor((2^92 - 1) / (span << s)),
// which spreads [0, span) over [0, 2^28) evenly (G is only monotone, not injective: ties inside a
// bin are ordered by the full hi' as before).  With on == 0, G is hi >> 36 and nothing changes.
// base and span come from the sampled minimum and maximum plus a margin; a key outside
// [base, base + span) raises the fallback flag.
struct KeyXform {
  uint64_t base;
  uint64_t span;
  uint64_t R;
  uint32_t s;
  uint32_t on;
};
__device__ __forceinline__ uint32_t g28(uint64_t hix, const KeyXform& x) {
  return x.on ? static_cast<uint32_t>(__umul64hi(hix, x.R)) : static_cast<uint32_t>(hix >> 36);
}
// the same with the choice made at compile time (pass and bucket kernels: no cost when off)
template <bool XF>
__device__ __forceinline__ uint32_t g28t(uint64_t hix, const KeyXform& x) {
  return XF ? static_cast<uint32_t>(__umul64hi(hix, x.R)) : static_cast<uint32_t>(hix >> 36);
}
static thread_local KeyXform t_xf = {0, 0, 0, 0, 0};  // of the sort call running on this thread

// The transform for n keys whose sample spans [mn, mx] (both in the encoded, unsigned order):
// the range widened by 1/64 on either side.  on == 0 when there is nothing to gain or the range
// holds fewer than n / 4 distinct leading words (runs of equal G beyond the repair limit).
static KeyXform make_key_xform(uint64_t mn, uint64_t mx, uint64_t n) {
  KeyXform xf = {0, 0, 0, 0, 0};
  if (mx <= mn) return xf;
  const uint64_t margin = (mx - mn) / 64 + 1;
  xf.base = mn > margin ? mn - margin : 0;
  const uint64_t top = mx < UINT64_MAX - margin ? mx + margin : UINT64_MAX;
  if (top - xf.base == UINT64_MAX || top - xf.base < n / 4) return xf;
  xf.span = top - xf.base + 1;
  xf.s = static_cast<uint32_t>(__builtin_clzll(xf.span));
  xf.R = static_cast<uint64_t>(((static_cast<unsigned __int128>(1) << 92) - 1) / (xf.span << xf.s));
  xf.on = 1;
  return xf;
}

// ---------------------------------------------------------------- encode + hist
// UNIFORM_CHECK: test every digit for being constant across the warp first (one atomic instead
// of 32 colliding ones on skewed keys).  Off when the caller already knows the digits spread
// (prefix mode, Philox keys): the four match.all per row are not free on sm_100a.
template <int KIND, bool HAS_LO, bool UNIFORM_CHECK>
__global__ void __launch_bounds__(512) encode_hist_kernel(const void* __restrict__ col,
                                                          uint32_t width, uint32_t n, bool desc,
                                                          uint64_t* __restrict__ hi_out,
                                                          uint16_t* __restrict__ lo_out, PassList pl,
                                                          uint32_t* __restrict__ ghist,
                                                          uint64_t seed, uint64_t subseq, KeyXform xf,
                                                          uint32_t* __restrict__ out_of_range) {
  __shared__ uint32_t sh[MAX_PASSES * RADIX];
  for (int i = threadIdx.x; i < pl.count * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t hi;
    uint32_t lo;
    if (KIND == KIND_PHILOX) {
      hi = philox64(seed, subseq, i);
      lo = 0;
    } else {
      load_key<KIND>(col, width, i, desc, hi, lo);
    }
    if (xf.on) {  // sub-range keys: store hi' (see KeyXform)
      uint64_t d0 = hi - xf.base;
      if (hi < xf.base || d0 >= xf.span) {
        atomicOr(out_of_range, 1u);
        d0 = hi < xf.base ? 0 : xf.span - 1;
      }
      hi = d0 << xf.s;
    }
    st_stream(hi_out + i, hi);
    if (HAS_LO) st_stream(lo_out + i, static_cast<uint16_t>(lo));
    const unsigned m = __activemask();
    const int leader = __ffs(m) - 1;
#pragma unroll
    for (int p = 0; p < MAX_PASSES; ++p) {
      if (p < pl.count) {
        uint32_t d = (HAS_LO && pl.from_lo[p]) ? (lo >> pl.shift[p]) & 255u
                                               : static_cast<uint32_t>(hi >> pl.shift[p]) & 255u;
        if (xf.on) d = (g28(hi, xf) >> pl.shift[p]) & 255u;  // shift = 12 / 20: the two bucket digits
        int same = 0;
        if (UNIFORM_CHECK) __match_all_sync(m, d, &same);
        if (same) {  // constant digit across the warp: one atomic instead of 32 colliding
          if ((int)lane_id() == leader) atomicAdd(&sh[p * RADIX + d], __popc(m));
        } else {
          atomicAdd(&sh[p * RADIX + d], 1u);
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < pl.count * RADIX; i += blockDim.x) {
    uint32_t c = sh[i];
    if (c) atomicAdd(&ghist[i], c);
  }
}

// Digit histograms of every `stride`-th key only: a cheap estimate of how well the leading
// digits separate the rows (the plan choice in sort_indices_impl), nothing is written back.
template <int KIND>
__global__ void __launch_bounds__(256) sample_hist_kernel(const void* __restrict__ col, uint32_t width,
                                                          uint32_t n, bool desc, uint32_t stride,
                                                          PassList pl, uint32_t* __restrict__ ghist,
                                                          unsigned long long* __restrict__ minmax, KeyXform xf) {
  __shared__ uint32_t sh[MAX_PASSES * RADIX];
  // minmax[0] = max(~hi) (i.e. ~min), minmax[1] = max(hi) over the sample; both start at 0
  __shared__ unsigned long long s_mm[2];
  if (threadIdx.x < 2) s_mm[threadIdx.x] = 0;
  for (int i = threadIdx.x; i < pl.count * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t ns = (n + stride - 1) / stride;
  unsigned long long t_inv_min = 0, t_max = 0;
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < ns; j += gridDim.x * blockDim.x) {
    uint64_t hi;
    uint32_t lo;
    load_key<KIND>(col, width, static_cast<uint64_t>(j) * stride, desc, hi, lo);
    t_inv_min = max(t_inv_min, static_cast<unsigned long long>(~hi));
    t_max = max(t_max, static_cast<unsigned long long>(hi));
    if (xf.on) {
      const uint64_t d0 = hi - xf.base;
      hi = ((hi < xf.base) ? 0 : (d0 >= xf.span ? xf.span - 1 : d0)) << xf.s;
    }
#pragma unroll
    for (int p = 0; p < MAX_PASSES; ++p) {
      if (p < pl.count) {
        uint32_t d = pl.from_lo[p] ? (lo >> pl.shift[p]) & 255u
                                   : static_cast<uint32_t>(hi >> pl.shift[p]) & 255u;
        if (xf.on) d = (g28(hi, xf) >> pl.shift[p]) & 255u;
        atomicAdd(&sh[p * RADIX + d], 1u);
      }
    }
  }
  if (minmax != nullptr) {
    atomicMax(&s_mm[0], t_inv_min);
    atomicMax(&s_mm[1], t_max);
  }
  __syncthreads();
  if (minmax != nullptr && threadIdx.x < 2) atomicMax(minmax + threadIdx.x, s_mm[threadIdx.x]);
  for (int i = threadIdx.x; i < pl.count * RADIX; i += blockDim.x) {
    const uint32_t c = sh[i];
    if (c) atomicAdd(&ghist[i], c);
  }
}

// One block of 256 threads per digit position.
__global__ void scan_hist_kernel(const uint32_t* __restrict__ ghist, uint32_t n,
                                 uint32_t* __restrict__ bin_base, uint32_t* __restrict__ skip,
                                 float* __restrict__ coll) {
  __shared__ uint32_t warp_tot[8];
  __shared__ float warp_sq[8];
  const int p = blockIdx.x, d = threadIdx.x;
  const uint32_t c = ghist[p * RADIX + d];
  uint32_t v = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if ((int)lane_id() >= o) v += t;
  }
  if (lane_id() == 31) warp_tot[d >> 5] = v;
  __syncthreads();
  uint32_t add = 0;
  for (int w = 0; w < (d >> 5); ++w) add += warp_tot[w];
  bin_base[p * RADIX + d] = v + add - c;
  const int any = __syncthreads_or(c == n);
  if (d == 0) skip[p] = any ? 1u : 0u;
  // probability that two random rows share this digit: sum_b (c_b / n)^2
  float f = static_cast<float>(c) / static_cast<float>(n);
  f *= f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) f += __shfl_xor_sync(0xffffffffu, f, o);
  if (lane_id() == 0) warp_sq[d >> 5] = f;
  __syncthreads();
  if (d == 0) {
    float t = 0;
    for (int w = 0; w < 8; ++w) t += warp_sq[w];
    coll[p] = t;
  }
}

// ---------------------------------------------------------------- prefix sort + run repair
// After sorting (stably) by the key bits >= prefix_shift only, rows whose prefixes are equal
// form runs in input order.  One thread per run head sorts its run by the full key with a
// stable insertion sort.  Runs longer than REPAIR_MAX_RUN raise *too_long (the caller then
// falls back to the full LSD sort).
constexpr int REPAIR_MAX_RUN = 16;

// Slow path of the repair: the run of equal prefixes that starts at row i (its head) is
// ordered by the full key with a stable insertion sort in local memory.
// Returns the run's length if rows were moved, else 0.
template <bool HAS_LO>
__device__ __noinline__ uint32_t repair_one_run(uint64_t* __restrict__ hi, uint16_t* __restrict__ lo,
                                                uint32_t* __restrict__ idx, uint32_t i, uint32_t n,
                                                int prefix_shift, uint32_t* __restrict__ too_long) {
  const uint64_t pfx = hi[i] >> prefix_shift;
  uint32_t len = 2;
  while (i + len < n && len <= REPAIR_MAX_RUN && (hi[i + len] >> prefix_shift) == pfx) ++len;
  if (len > REPAIR_MAX_RUN) {
    atomicOr(too_long, 1u);
    return 0;
  }
  uint64_t h[REPAIR_MAX_RUN];
  uint32_t l[REPAIR_MAX_RUN], x[REPAIR_MAX_RUN];
  for (uint32_t k = 0; k < len; ++k) {
    h[k] = hi[i + k];
    l[k] = HAS_LO ? lo[i + k] : 0;
    x[k] = idx[i + k];
  }
  bool moved = false;
  for (uint32_t a = 1; a < len; ++a) {  // stable: only strictly smaller keys move left
    const uint64_t kh = h[a];
    const uint32_t kl = l[a], kx = x[a];
    uint32_t b = a;
    while (b > 0 && (h[b - 1] > kh || (h[b - 1] == kh && l[b - 1] > kl))) {
      h[b] = h[b - 1];
      l[b] = l[b - 1];
      x[b] = x[b - 1];
      --b;
      moved = true;
    }
    h[b] = kh;
    l[b] = kl;
    x[b] = kx;
  }
  if (!moved) return 0;
  for (uint32_t k = 0; k < len; ++k) {
    hi[i + k] = h[k];
    if (HAS_LO) lo[i + k] = static_cast<uint16_t>(l[k]);
    idx[i + k] = x[k];
  }
  return len;
}

// Each thread classifies REPAIR_VEC consecutive rows from two 16-byte loads (plus its two
// neighbours) and calls the slow path only for rows that head a run of two or more equal
// prefixes.  Runs never overlap, and only the thread that holds a run's head writes it, so
// the in-place update is race free; a neighbouring thread may read a prefix while the run
// is being rewritten, but prefixes of all members of a run are equal by definition.
constexpr int REPAIR_VEC = 4;
template <bool HAS_LO>
__global__ void __launch_bounds__(256)
    repair_runs_kernel(uint64_t* __restrict__ hi, uint16_t* __restrict__ lo, uint32_t* __restrict__ idx,
                       uint32_t n, int prefix_shift, uint32_t* __restrict__ too_long) {
  const uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) * REPAIR_VEC;
  if (i0 >= n) return;
  uint64_t p[REPAIR_VEC + 2];  // p[k + 1] = prefix of row i0 + k
  const bool whole = i0 + REPAIR_VEC <= n;
  if (whole) {
    const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(hi + i0);
    const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(hi + i0 + 2);
    p[1] = a.x >> prefix_shift;
    p[2] = a.y >> prefix_shift;
    p[3] = b.x >> prefix_shift;
    p[4] = b.y >> prefix_shift;
  } else {
#pragma unroll
    for (int k = 0; k < REPAIR_VEC; ++k) p[k + 1] = i0 + k < n ? hi[i0 + k] >> prefix_shift : 0;
  }
  const bool has_prev = i0 > 0, has_next = i0 + REPAIR_VEC < n;
  p[0] = has_prev ? hi[i0 - 1] >> prefix_shift : 0;
  p[REPAIR_VEC + 1] = has_next ? hi[i0 + REPAIR_VEC] >> prefix_shift : 0;
  uint32_t heads = 0;
#pragma unroll
  for (int k = 0; k < REPAIR_VEC; ++k) {
    const uint32_t i = i0 + k;
    const bool same_prev = (k > 0 || has_prev) && p[k] == p[k + 1];
    const bool same_next = (k + 1 < REPAIR_VEC ? (whole || i + 1 < n) : has_next) && p[k + 1] == p[k + 2];
    if (i < n && !same_prev && same_next) heads |= 1u << k;
  }
  while (heads) {  // one divergent region per warp instead of one per k
    const int k = __ffs(heads) - 1;
    heads &= heads - 1;
    repair_one_run<HAS_LO>(hi, lo, idx, i0 + k, n, prefix_shift, too_long);
  }
}

__global__ void iota_kernel(uint32_t* out, uint32_t n) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = i;
}

// ---------------------------------------------------------------- onesweep pass
template <bool HAS_LO, bool DIGIT_LO>
__device__ __forceinline__ uint32_t digit_of(uint64_t hi, uint32_t lo, int shift) {
  return DIGIT_LO && HAS_LO ? (lo >> shift) & 255u : static_cast<uint32_t>(hi >> shift) & 255u;
}
// the same with the sub-range transform: shift then counts bits of G (12 / 20)
template <bool HAS_LO, bool DIGIT_LO, bool XF>
__device__ __forceinline__ uint32_t digit_x(uint64_t hi, uint32_t lo, int shift, const KeyXform& xf) {
  return XF ? (g28t<true>(hi, xf) >> shift) & 255u : digit_of<HAS_LO, DIGIT_LO>(hi, lo, shift);
}

template <int THREADS, int ITEMS, bool HAS_LO>
struct PassSmem {
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr int WARPS = THREADS / 32;
  static constexpr int ITEM_BYTES = 8 + (HAS_LO ? 2 : 0) + 4;
  static constexpr int EXCH = TILE * ITEM_BYTES;
  static constexpr int HIST = WARPS * RADIX * 4;
  static constexpr int BYTES = (EXCH > HIST ? EXCH : HIST);
};

// helpers of the pass kernel (onesweep_v3.cuh)
__device__ __forceinline__ uint32_t ballot_match8(uint32_t d) {
  uint32_t peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    const bool bit = (d >> b) & 1u;
    const uint32_t v = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? v : ~v;
  }
  return peers;
}
__device__ __forceinline__ uint32_t ld_relaxed_gpu(const volatile uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_gpu(volatile uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

}  // namespace ds
#include "onesweep_v3.cuh"
namespace ds {
__device__ __forceinline__ uint64_t bswap64(uint64_t v);
}
#include "bucket_sort.cuh"
namespace ds {

// ---------------------------------------------------------------- decode
template <int KIND>
__global__ void decode_kernel(const uint64_t* __restrict__ hi, const uint16_t* __restrict__ lo,
                              uint32_t width, uint32_t n, bool desc, void* __restrict__ out) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t h = hi[i];
  if (KIND == KIND_I64) {
    static_cast<int64_t*>(out)[i] = dec_i64(h, desc);
  } else if (KIND == KIND_U64) {
    static_cast<uint64_t*>(out)[i] = desc ? ~h : h;
  } else {
    uint32_t l = width > 8 ? lo[i] : 0;
    if (desc) {
      h = ~h;
      l = ~l;
    }
    uint8_t* p = static_cast<uint8_t*>(out) + static_cast<uint64_t>(i) * width;
#pragma unroll
    for (int b = 0; b < 8; ++b)
      if (b < (int)width) p[b] = static_cast<uint8_t>(h >> (56 - 8 * b));
    if (width == 9) p[8] = static_cast<uint8_t>(l);
    if (width == 10) {
      p[8] = static_cast<uint8_t>(l >> 8);
      p[9] = static_cast<uint8_t>(l);
    }
  }
}

// fixed_size_binary decode with coalesced 16-byte stores: DECODE_ROWS rows are laid out in
// shared memory exactly as they appear in the output buffer, then copied out as uint4.
// Each thread decodes DECODE_ROWS / 256 rows so that several loads are in flight.
constexpr int DECODE_ROWS = 1024;
__device__ __forceinline__ uint64_t bswap64(uint64_t v) {
  const uint32_t lo32 = static_cast<uint32_t>(v), hi32 = static_cast<uint32_t>(v >> 32);
  return (static_cast<uint64_t>(__byte_perm(lo32, 0, 0x0123)) << 32) | __byte_perm(hi32, 0, 0x0123);
}
template <bool W10>
__global__ void __launch_bounds__(256)
    decode_fsb_staged_kernel(const uint64_t* __restrict__ hi, const uint16_t* __restrict__ lo,
                             uint32_t width, uint32_t n, bool desc, uint8_t* __restrict__ out) {
  constexpr int PER = DECODE_ROWS / 256;
  __shared__ __align__(16) uint8_t buf[DECODE_ROWS * DS_MAX_KEY_BYTES];
  const uint32_t row0 = blockIdx.x * static_cast<uint32_t>(DECODE_ROWS);
  uint64_t h[PER];
  uint32_t l[PER];
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    const uint32_t i = row0 + k * 256 + threadIdx.x;
    h[k] = i < n ? ld_stream(hi + i) : 0;
    l[k] = (i < n && width > 8) ? ld_stream(lo + i) : 0;
  }
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    if (desc) {
      h[k] = ~h[k];
      l[k] = ~l[k];
    }
    const uint32_t r = k * 256 + threadIdx.x;
    if (W10) {  // rows start on 2-byte boundaries: five 16-bit stores
      const uint64_t be = bswap64(h[k]);
      uint16_t* q = reinterpret_cast<uint16_t*>(buf + r * 10);
      q[0] = static_cast<uint16_t>(be);
      q[1] = static_cast<uint16_t>(be >> 16);
      q[2] = static_cast<uint16_t>(be >> 32);
      q[3] = static_cast<uint16_t>(be >> 48);
      q[4] = static_cast<uint16_t>(((l[k] >> 8) & 0xffu) | ((l[k] & 0xffu) << 8));
    } else {
      uint8_t* q = buf + r * width;
#pragma unroll
      for (int b = 0; b < 8; ++b)
        if (b < (int)width) q[b] = static_cast<uint8_t>(h[k] >> (56 - 8 * b));
      if (width == 9) q[8] = static_cast<uint8_t>(l[k]);
      if (width == 10) {
        q[8] = static_cast<uint8_t>(l[k] >> 8);
        q[9] = static_cast<uint8_t>(l[k]);
      }
    }
  }
  __syncthreads();
  const uint32_t nbytes = min(static_cast<uint32_t>(DECODE_ROWS), n - row0) * width;
  uint8_t* dst = out + static_cast<uint64_t>(row0) * width;  // 16-byte aligned: 1024*width % 16 == 0
  const uint32_t vec = nbytes >> 4;
  for (uint32_t v = threadIdx.x; v < vec; v += 256)
    st_stream(reinterpret_cast<uint4*>(dst) + v, reinterpret_cast<const uint4*>(buf)[v]);
  for (uint32_t o = (vec << 4) + threadIdx.x; o < nbytes; o += 256) dst[o] = buf[o];
}

// Prefix mode, 10-byte keys: the decode and the search for runs of equal prefixes read the
// same sorted hi/lo stream, so one kernel does both -- it writes the key column as if no run
// needed repair and appends the position of every run head to head_list (one global atomic
// per block).  repair_listed_runs_kernel then visits only those heads (about 1 % of the rows),
// orders each run by the full key and rewrites the run's 10-byte keys in the output.
// Thread t of a block owns rows row0 + 4t .. 4t+3: two 16-byte loads of hi, one 8-byte load of
// lo, 40 contiguous output bytes staged in shared memory as five 8-byte words.
__global__ void __launch_bounds__(256)
    decode10_find_runs_kernel(const uint64_t* __restrict__ hi, const uint16_t* __restrict__ lo, uint32_t n,
                              bool desc, int prefix_shift, uint8_t* __restrict__ out,
                              uint32_t* __restrict__ head_list, uint32_t* __restrict__ head_count) {
  __shared__ __align__(16) uint8_t buf[DECODE_ROWS * 10];
  __shared__ uint32_t s_warp[8];
  __shared__ uint32_t s_base;
  const uint32_t row0 = blockIdx.x * static_cast<uint32_t>(DECODE_ROWS);
  const uint32_t i0 = row0 + threadIdx.x * 4;
  uint64_t h[4] = {0, 0, 0, 0};
  uint32_t l[4] = {0, 0, 0, 0};
  const bool whole = i0 + 4 <= n;
  if (whole) {
    const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(hi + i0);
    const ulonglong2 b = *reinterpret_cast<const ulonglong2*>(hi + i0 + 2);
    const ushort4 c = *reinterpret_cast<const ushort4*>(lo + i0);
    h[0] = a.x; h[1] = a.y; h[2] = b.x; h[3] = b.y;
    l[0] = c.x; l[1] = c.y; l[2] = c.z; l[3] = c.w;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (i0 + k < n) {
        h[k] = hi[i0 + k];
        l[k] = lo[i0 + k];
      }
  }
  // ---- run heads among my rows (same rule as repair_runs_kernel)
  uint32_t heads = 0;
  if (i0 < n) {
    uint64_t p[6];
#pragma unroll
    for (int k = 0; k < 4; ++k) p[k + 1] = h[k] >> prefix_shift;
    const bool has_prev = i0 > 0, has_next = i0 + 4 < n;
    p[0] = has_prev ? hi[i0 - 1] >> prefix_shift : 0;
    p[5] = has_next ? hi[i0 + 4] >> prefix_shift : 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t i = i0 + k;
      const bool same_prev = (k > 0 || has_prev) && p[k] == p[k + 1];
      const bool same_next = (k + 1 < 4 ? (whole || i + 1 < n) : has_next) && p[k + 1] == p[k + 2];
      if (i < n && !same_prev && same_next) heads |= 1u << k;
    }
  }
  // ---- decode into the staging buffer: 4 rows = 40 bytes = 5 little-endian words
  {
    uint64_t B[4], L[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint64_t hh = desc ? ~h[k] : h[k];
      const uint32_t ll = (desc ? ~l[k] : l[k]) & 0xffffu;
      B[k] = bswap64(hh);
      L[k] = ((ll >> 8) | (ll << 8)) & 0xffffu;
    }
    uint64_t* q = reinterpret_cast<uint64_t*>(buf + threadIdx.x * 40);
    q[0] = B[0];
    q[1] = L[0] | (B[1] << 16);
    q[2] = (B[1] >> 48) | (L[1] << 16) | (B[2] << 32);
    q[3] = (B[2] >> 32) | (L[2] << 32) | (B[3] << 48);
    q[4] = (B[3] >> 16) | (L[3] << 48);
  }
  // ---- slots for my heads: warp scan, block offsets, one global atomic per block
  const uint32_t cnt = __popc(heads);
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane_id() >= o) incl += v;
  }
  if (lane_id() == 31) s_warp[threadIdx.x >> 5] = incl;
  __syncthreads();  // staging buffer and warp totals complete
  if (threadIdx.x == 0) {
    uint32_t tot = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += s_warp[w];
    s_base = tot ? atomicAdd(head_count, tot) : 0u;
  }
  const uint32_t nbytes = min(static_cast<uint32_t>(DECODE_ROWS), n - row0) * 10u;
  uint8_t* dst = out + static_cast<uint64_t>(row0) * 10u;
  const uint32_t vec = nbytes >> 4;
  for (uint32_t v = threadIdx.x; v < vec; v += 256)
    st_stream(reinterpret_cast<uint4*>(dst) + v, reinterpret_cast<const uint4*>(buf)[v]);
  for (uint32_t o = (vec << 4) + threadIdx.x; o < nbytes; o += 256) dst[o] = buf[o];
  __syncthreads();  // s_base
  if (heads) {
    uint32_t slot = s_base + incl - cnt;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) slot += s_warp[w];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (heads >> k & 1u) head_list[slot++] = i0 + k;
  }
}

__global__ void __launch_bounds__(256)
    repair_listed_runs_kernel(uint64_t* __restrict__ hi, uint16_t* __restrict__ lo, uint32_t* __restrict__ idx,
                              uint32_t n, int prefix_shift, const uint32_t* __restrict__ head_list,
                              const uint32_t* __restrict__ head_count, bool desc, uint8_t* __restrict__ out,
                              uint32_t* __restrict__ too_long) {
  const uint32_t cnt = *head_count;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < cnt; j += stride) {
    const uint32_t i = head_list[j];
    const uint32_t len = repair_one_run<true>(hi, lo, idx, i, n, prefix_shift, too_long);
    for (uint32_t r = 0; r < len; ++r) {  // rewrite the keys of the rows that moved
      const uint64_t hh = desc ? ~hi[i + r] : hi[i + r];
      const uint32_t ll = (desc ? ~static_cast<uint32_t>(lo[i + r]) : lo[i + r]) & 0xffffu;
      const uint64_t be = bswap64(hh);
      uint16_t* q = reinterpret_cast<uint16_t*>(out + static_cast<uint64_t>(i + r) * 10u);
      q[0] = static_cast<uint16_t>(be);
      q[1] = static_cast<uint16_t>(be >> 16);
      q[2] = static_cast<uint16_t>(be >> 32);
      q[3] = static_cast<uint16_t>(be >> 48);
      q[4] = static_cast<uint16_t>((ll >> 8) | ((ll & 0xffu) << 8));
    }
  }
}

// ---------------------------------------------------------------- host driver
static std::atomic<int> g_hybrid_mode{1};  // ds_debug_set_hybrid: 0 never, 1 when the buckets fill a CTA, 2 whatever n (tests)
static std::atomic<bool> g_no_subrange{false};  // ds_debug_set_subrange(0): no hybrid plan on a sub-range of the key space
static std::atomic<int> g_last_plan{-1};        // ds_debug_last_sort_plan
static std::atomic<bool> g_no_fused_decode{false};  // ds_debug_set_fused_decode(0): separate repair + decode kernels
static std::atomic<int> g_cfg{18};  // tuning knob (ds_debug_set_radix_config); 18 = v3 table match <384 x 12>, best measured

// one bit per device; relaxed atomics: concurrent first launches set the same attribute twice
using DoneMask = std::atomic<unsigned long long>;
template <typename K>
static cudaError_t set_smem_attr(K k, int bytes, DoneMask* done_mask) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(done_mask->load(std::memory_order_relaxed) >> (dev & 63) & 1ull)) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    done_mask->fetch_or(1ull << (dev & 63), std::memory_order_relaxed);
  }
  return cudaSuccess;
}

template <int VARIANT, int THREADS, int ITEMS, int MIN_BLOCKS, bool HAS_LO, bool DIGIT_LO, bool GEN_IDX>
static cudaError_t launch_pass_t(const uint64_t* hi_in, const uint16_t* lo_in,
                                 const uint32_t* idx_in, uint64_t* hi_out, uint16_t* lo_out,
                                 uint32_t* idx_out, uint32_t n, int shift, const uint32_t* bin_base,
                                 uint32_t* lookback, uint32_t* tile_counter, cudaStream_t s) {
  using S = PassSmem<THREADS, ITEMS, HAS_LO>;
  static DoneMask done_a{0}, done_b{0};  // per instantiation, one bit per device
  const uint32_t tiles = (n + S::TILE - 1) / S::TILE;
  cudaError_t e;
  LaunchTimer timer(TIMED_PASS, s, static_cast<double>(n) * (HAS_LO ? 28.0 : 24.0));
  static_assert(VARIANT >= 4 && VARIANT <= 6, "the first two generations of the pass kernel are retired");
  {
    constexpr bool TABLE = VARIANT == 4 || VARIANT == 5;
    constexpr bool PACKED = VARIANT == 4 || VARIANT == 6;
    using S3 = PassSmemV3<THREADS, ITEMS, PACKED, HAS_LO>;
    // full tiles first, then (stream-ordered) the ragged last tile as tile id `full_tiles`
    const uint32_t full_tiles = n / S::TILE;
    auto launch = [&](auto kf, auto kt, DoneMask* da, DoneMask* db) -> cudaError_t {
      cudaError_t e2;
      if ((e2 = set_smem_attr(kf, S3::BYTES, da)) != cudaSuccess) return e2;
      if ((e2 = set_smem_attr(kt, S3::BYTES, db)) != cudaSuccess) return e2;
      if (full_tiles) {
        ds::note_launch();
        kf<<<full_tiles, THREADS, S3::BYTES, s>>>(hi_in, lo_in, idx_in, hi_out, lo_out, idx_out, n,
                                                  shift, bin_base, lookback, tile_counter, t_xf);
      }
      if (tiles > full_tiles) {
        ds::note_launch();
        kt<<<1, THREADS, S3::BYTES, s>>>(hi_in, lo_in, idx_in, hi_out, lo_out, idx_out, n, shift,
                                         bin_base, lookback, tile_counter, t_xf);
      }
      return cudaSuccess;
    };
    if (t_xf.on) {  // sub-range hybrid plan: digits of G, always taken from `hi`
      if constexpr (!DIGIT_LO) {
        static DoneMask done_c{0}, done_d{0};
        e = launch(onesweep_pass_kernel_v3<THREADS, ITEMS, MIN_BLOCKS, HAS_LO, false, GEN_IDX, true, TABLE, PACKED, true>,
                   onesweep_pass_kernel_v3<THREADS, ITEMS, MIN_BLOCKS, HAS_LO, false, GEN_IDX, false, TABLE, PACKED, true>,
                   &done_c, &done_d);
      } else {
        e = cudaErrorInvalidValue;
      }
    } else {
      e = launch(onesweep_pass_kernel_v3<THREADS, ITEMS, MIN_BLOCKS, HAS_LO, DIGIT_LO, GEN_IDX, true, TABLE, PACKED, false>,
                 onesweep_pass_kernel_v3<THREADS, ITEMS, MIN_BLOCKS, HAS_LO, DIGIT_LO, GEN_IDX, false, TABLE, PACKED, false>,
                 &done_a, &done_b);
    }
    if (e != cudaSuccess) return e;
  }
  timer.stop();
  return cudaGetLastError();
}

template <int VARIANT, int THREADS, int ITEMS, int MIN_BLOCKS>
static cudaError_t launch_pass_cfg(bool has_lo, bool digit_lo, bool gen_idx, const uint64_t* hi_in,
                                   const uint16_t* lo_in, const uint32_t* idx_in, uint64_t* hi_out,
                                   uint16_t* lo_out, uint32_t* idx_out, uint32_t n, int shift,
                                   const uint32_t* bin_base, uint32_t* lookback,
                                   uint32_t* tile_counter, cudaStream_t s) {
#define DS_PASS(HL, DL, GI)                                                                      \
  return launch_pass_t<VARIANT, THREADS, ITEMS, MIN_BLOCKS, HL, DL, GI>(hi_in, lo_in, idx_in, hi_out,     \
                                                               lo_out, idx_out, n, shift,        \
                                                               bin_base, lookback, tile_counter, s)
  if (has_lo) {
    if (digit_lo) {
      if (gen_idx) DS_PASS(true, true, true);
      DS_PASS(true, true, false);
    }
    if (gen_idx) DS_PASS(true, false, true);
    DS_PASS(true, false, false);
  }
  if (gen_idx) DS_PASS(false, false, true);
  DS_PASS(false, false, false);
#undef DS_PASS
}

// (variant, threads, items, min blocks per SM) of onesweep_pass_kernel_v3: variant 4 = match
// table + packed exchange slots, 5 = match table (the default, cfg 18), 6 = ballots + packed.
// The first two generations of the kernel (ids 0..11) were retired after round 1; their
// measurements are in profiles/r01_pass_kernel_v2_summary.txt.
#define DS_PASS_CONFIGS(X) \
  X(17, 4, 384, 12, 2) X(18, 5, 384, 12, 2) X(19, 6, 384, 12, 2) \
  X(20, 4, 512, 8, 2) X(21, 5, 256, 12, 3) X(25, 5, 512, 8, 2) \
  X(28, 5, 256, 16, 2)
constexpr int NUM_CFGS = 29;  // ids are sparse: unknown ids are rejected

static bool cfg_known(int cfg) {
  switch (cfg) {
#define X(id, v, t, i, m) case id:
    DS_PASS_CONFIGS(X)
#undef X
    return true;
    default: return false;
  }
}

static int cfg_tile(int cfg) {
  switch (cfg) {
#define X(id, v, t, i, m) case id: return t * i;
    DS_PASS_CONFIGS(X)
#undef X
    default: return 256 * 16;
  }
}

constexpr uint32_t min_cfg_tile() {
  uint32_t m = 0xffffffffu;
#define X(id, v, t, i, m_) m = (t) * (i) < m ? (t) * (i) : m;
  DS_PASS_CONFIGS(X)
#undef X
  return m;
}
constexpr uint32_t MIN_CFG_TILE = min_cfg_tile();

static cudaError_t launch_pass(int cfg, bool has_lo, bool digit_lo, bool gen_idx,
                               const uint64_t* hi_in, const uint16_t* lo_in, const uint32_t* idx_in,
                               uint64_t* hi_out, uint16_t* lo_out, uint32_t* idx_out, uint32_t n,
                               int shift, const uint32_t* bin_base, uint32_t* lookback,
                               uint32_t* tile_counter, cudaStream_t s) {
  switch (cfg) {
#define X(id, v, t, i, m)                                                                        \
  case id:                                                                                       \
    return launch_pass_cfg<v, t, i, m>(has_lo, digit_lo, gen_idx, hi_in, lo_in, idx_in, hi_out,  \
                                       lo_out, idx_out, n, shift, bin_base, lookback,            \
                                       tile_counter, s);
    DS_PASS_CONFIGS(X)
#undef X
    default: return cudaErrorInvalidValue;
  }
}

static PassList make_passes(int kind, uint32_t width) {
  PassList pl{};
  if (kind != KIND_FSB) {
    pl.count = 8;
    for (int p = 0; p < 8; ++p) {
      pl.shift[p] = 8 * p;
      pl.from_lo[p] = 0;
    }
    return pl;
  }
  int c = 0;
  if (width > 8) {
    for (uint32_t b = 0; b < width - 8; ++b) {
      pl.shift[c] = 8 * b;
      pl.from_lo[c] = 1;
      ++c;
    }
  }
  const int hi_bytes = width < 8 ? (int)width : 8;
  for (int b = 0; b < hi_bytes; ++b) {
    pl.shift[c] = 64 - 8 * hi_bytes + 8 * b;
    pl.from_lo[c] = 0;
    ++c;
  }
  pl.count = c;
  return pl;
}

template <int KIND>
static cudaError_t launch_encode(bool has_lo, bool uniform_check, const void* col, uint32_t width, uint32_t n,
                                 bool desc, uint64_t* hi, uint16_t* lo, const PassList& pl, uint32_t* ghist,
                                 cudaStream_t s, uint64_t seed = 0, uint64_t subseq = 0,
                                 uint32_t* out_of_range = nullptr) {
  const int threads = 512;
  int blocks = sm_count() * 4;
  const int need = (n + threads - 1) / threads;
  if (blocks > need) blocks = need;
  note_launch();
  if (has_lo && uniform_check)
    encode_hist_kernel<KIND, true, true><<<blocks, threads, 0, s>>>(col, width, n, desc, hi, lo, pl, ghist, seed, subseq, t_xf, out_of_range);
  else if (has_lo)
    encode_hist_kernel<KIND, true, false><<<blocks, threads, 0, s>>>(col, width, n, desc, hi, lo, pl, ghist, seed, subseq, t_xf, out_of_range);
  else if (uniform_check)
    encode_hist_kernel<KIND, false, true><<<blocks, threads, 0, s>>>(col, width, n, desc, hi, lo, pl, ghist, seed, subseq, t_xf, out_of_range);
  else
    encode_hist_kernel<KIND, false, false><<<blocks, threads, 0, s>>>(col, width, n, desc, hi, lo, pl, ghist, seed, subseq, t_xf, out_of_range);
  return cudaGetLastError();
}

ds_status gather_rows_impl(const void* src, uint32_t row_bytes, uint64_t src_rows,
                           const uint32_t* idx, uint64_t n, void* dst, cudaStream_t s);  // gather.cu

static ds_status sort_indices_impl(int kind, const void* col, uint32_t width, uint64_t n64,
                                   int32_t descending, uint32_t* idx_out, void* sorted_out,
                                   int32_t options, void* ws, size_t* ws_bytes, void* stream,
                                   const char* fn, uint64_t seed, uint64_t subseq);

static ds_status sort_indices_impl(int kind, const void* col, uint32_t width, uint64_t n64,
                                   int32_t descending, uint32_t* idx_out, void* sorted_out,
                                   int32_t options, void* ws, size_t* ws_bytes, void* stream,
                                   const char* fn, uint64_t seed, uint64_t subseq) {
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: n=%llu exceeds DS_MAX_ROWS", fn, (unsigned long long)n64);
    return DS_ERR_TOO_LARGE;
  }
  if (kind == KIND_FSB && (width < 1 || width > DS_MAX_KEY_BYTES)) {
    set_error("%s: fixed_size_binary key width %u unsupported (1..%d)", fn, width,
              DS_MAX_KEY_BYTES);
    return DS_ERR_UNSUPPORTED;
  }
  const uint32_t n = static_cast<uint32_t>(n64);
  const bool has_lo = kind == KIND_FSB && width > 8;
  const PassList pl = make_passes(kind, width);
  t_xf = KeyXform{0, 0, 0, 0, 0};  // set again below if this call takes the sub-range hybrid plan
  const int cfg = g_cfg.load(std::memory_order_relaxed);
  const uint32_t tiles = (n + cfg_tile(cfg) - 1) / cfg_tile(cfg);

  Workspace w(ws);
  uint64_t* hi[2] = {w.take<uint64_t>(n), w.take<uint64_t>(n)};
  uint16_t* lo[2] = {nullptr, nullptr};
  if (has_lo) {
    lo[0] = w.take<uint16_t>(n);
    lo[1] = w.take<uint16_t>(n);
  }
  uint32_t* idx_alt = w.take<uint32_t>(n);
  uint32_t* ctrl = w.take<uint32_t>(MAX_PASSES * RADIX * 2 + 4 * MAX_PASSES);
  // sized for the smallest tile of any pass configuration: the size query and the run may see
  // different ds_debug_set_radix_config settings
  uint32_t* lookback = w.take<uint32_t>(static_cast<size_t>((n + MIN_CFG_TILE - 1) / MIN_CFG_TILE) * RADIX);
  uint32_t* bucket_starts = w.take<uint32_t>(LS_BUCKETS + 1);  // hybrid plan only (257 KB)
  ds_status st;
  if (ws_resolve(w, ws, ws_bytes, fn, &st)) return st;
  if (n == 0) return DS_OK;
  if ((col == nullptr && kind != KIND_PHILOX) || idx_out == nullptr) {
    set_error("%s: col and idx_out must not be NULL", fn);
    return DS_ERR_INVALID_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint32_t* ghist = ctrl;
  uint32_t* bin_base = ctrl + MAX_PASSES * RADIX;
  uint32_t* skip = ctrl + 2 * MAX_PASSES * RADIX;
  uint32_t* tile_counters = skip + MAX_PASSES;
  float* coll = reinterpret_cast<float*>(tile_counters + MAX_PASSES);  // directly after skip..: read back together
  uint32_t* too_long = tile_counters + 2 * MAX_PASSES;
  const bool desc = descending != 0;

  const size_t ctrl_bytes = sizeof(uint32_t) * (MAX_PASSES * RADIX * 2 + 4 * MAX_PASSES);
  DS_CUDA(cudaMemsetAsync(ctrl, 0, ctrl_bytes, s));
  cudaError_t e;

  // ---- plan.  Prefix sort + run repair: if the leading m digits alone already separate
  // almost all rows (expected colliding pairs <= 2% of n, digits taken as independent), run
  // only those m passes and let repair_runs_kernel order the few short runs of equal
  // prefixes by the full key.  The decision uses digit histograms of a ~1M-key sample (or
  // the known uniformity of Philox keys); it is only a heuristic: the repair kernel VERIFIES
  // every run, and a run it cannot handle sends the call back to the all-passes plan.
  PassList run = pl;  // the digit positions that get exact histograms (and passes)
  int prefix_shift = -1;
  bool hybrid = false;
  if (!(options & (DS_SORT_NO_PASS_SKIP | DS_SORT_NO_PREFIX_REPAIR)) && n >= (1u << 16) &&
      pl.count >= 3) {
    // read back in one copy: coll[MAX_PASSES], too_long, head_count, then the sampled min / max
    struct SampleReadback {
      float coll[MAX_PASSES];
      uint32_t flags[2];
      unsigned long long inv_min, max;
    } rb;
    static_assert(sizeof(SampleReadback) == 4 * (MAX_PASSES + 2) + 16, "layout of ctrl behind coll");
    float* const coll_h = rb.coll;
    unsigned long long* const minmax = reinterpret_cast<unsigned long long*>(too_long + 2);
    const uint32_t stride = n > (1u << 20) ? n >> 20 : 1;
    const uint32_t ns = (n + stride - 1) / stride;
    const int sample_blocks = static_cast<int>((ns + 255) / 256 < 1184u ? (ns + 255) / 256 : 1184u);
    auto sample = [&](const PassList& digits, unsigned long long* mm, const KeyXform& xf) {
      ds::note_launch();
      switch (kind) {
        case KIND_I64: sample_hist_kernel<KIND_I64><<<sample_blocks, 256, 0, s>>>(col, width, n, desc, stride, digits, ghist, mm, xf); break;
        case KIND_U64: sample_hist_kernel<KIND_U64><<<sample_blocks, 256, 0, s>>>(col, width, n, desc, stride, digits, ghist, mm, xf); break;
        case KIND_F64: sample_hist_kernel<KIND_F64><<<sample_blocks, 256, 0, s>>>(col, width, n, desc, stride, digits, ghist, mm, xf); break;
        default: sample_hist_kernel<KIND_FSB><<<sample_blocks, 256, 0, s>>>(col, width, n, desc, stride, digits, ghist, mm, xf); break;
      }
      ds::note_launch(); scan_hist_kernel<<<digits.count, RADIX, 0, s>>>(ghist, ns, bin_base, skip, coll);
    };
    if (kind == KIND_PHILOX) {
      for (int p = 0; p < pl.count; ++p) coll_h[p] = 1.0f / 256.0f;
    } else {
      sample(pl, minmax, KeyXform{0, 0, 0, 0, 0});
      DS_LAUNCH_CHECK();
      DS_CUDA(cudaMemcpyAsync(&rb, coll, sizeof(rb), cudaMemcpyDeviceToHost, s));
      DS_CUDA(cudaMemsetAsync(ctrl, 0, ctrl_bytes, s));
      DS_CUDA(cudaStreamSynchronize(s));  // host sync: 64 bytes
    }
    // hybrid plan (bucket_sort.cuh): 10-byte keys whose two leading bytes spread evenly and
    // whose 65536 buckets would fill a CTA reasonably (400..1650 rows on average)
    const bool fsb10 = kind == KIND_FSB && width == 10 && sorted_out != nullptr &&
                       (reinterpret_cast<uintptr_t>(sorted_out) & 15) == 0;
    const bool key64 = kind != KIND_FSB;  // i64 / u64 / f64 / Philox: the whole key is `hi`
    const int hybrid_mode = g_hybrid_mode.load(std::memory_order_relaxed);
    if ((fsb10 || key64) && hybrid_mode != 0 && !(options & DS_SORT_NO_HYBRID) &&
        (n >= 400u * LS_BUCKETS || hybrid_mode == 2) && n <= 1650u * LS_BUCKETS) {
      const float lim = 1.25f / 256.0f;
      hybrid = coll_h[pl.count - 1] <= lim && coll_h[pl.count - 2] <= lim;
      if (!hybrid && kind != KIND_PHILOX && !g_no_subrange.load(std::memory_order_relaxed)) {
        // The leading bytes do not spread: the keys may still fill a SUB-RANGE of the key space
        // evenly (see KeyXform).  Take the range from the sampled extremes plus 1/64 on either
        // side, and look at the digits of G over the same sample before committing to it.
        const KeyXform xf = make_key_xform(~rb.inv_min, rb.max, n);
        if (xf.on) {
          PassList g;  // bits [4,12), [12,20), [20,28) of G
          g.count = 3;
          for (int k = 0; k < 3; ++k) {
            g.shift[k] = 4 + 8 * k;
            g.from_lo[k] = 0;
          }
          float c[3];
          sample(g, nullptr, xf);
          DS_LAUNCH_CHECK();
          DS_CUDA(cudaMemcpyAsync(c, coll, sizeof(c), cudaMemcpyDeviceToHost, s));
          DS_CUDA(cudaMemsetAsync(ctrl, 0, ctrl_bytes, s));
          DS_CUDA(cudaStreamSynchronize(s));  // host sync: 12 bytes (sub-range candidates only)
          // expected pairs of rows with equal G (28 bits): the same head-room as the
          // prefix_shift >= 32 rule of the plain hybrid plan below
          const double pairs28 = 0.5 * static_cast<double>(n) * static_cast<double>(n) * c[0] * c[1] * c[2] / 16.0;
          if (c[1] <= lim && c[2] <= lim && pairs28 <= 0.32 * static_cast<double>(n)) {
            hybrid = true;
            t_xf = xf;
          }
        }
      }
    }
    double pairs = 0.5 * static_cast<double>(n) * static_cast<double>(n);
    for (int m = 1; m <= pl.count - 2; ++m) {
      const int p = pl.count - m;
      if (pl.from_lo[p]) break;  // the prefix must lie inside `hi`
      pairs *= static_cast<double>(coll_h[p]);
      if (pairs <= 0.02 * static_cast<double>(n)) {
        prefix_shift = pl.shift[p];
        run.count = m;
        for (int k = 0; k < m; ++k) {
          run.shift[k] = pl.shift[p + k];
          run.from_lo[k] = pl.from_lo[p + k];
        }
        break;
      }
    }
  }

  // the hybrid plan needs the prefix plan's guarantee (few ties on the leading 32 bits)
  hybrid = hybrid && (t_xf.on || prefix_shift >= 32);
  if (hybrid) {  // only the two leading digits run as global passes
    run.count = 2;
    for (int k = 0; k < 2; ++k) {
      run.shift[k] = t_xf.on ? 12 + 8 * k : pl.shift[pl.count - 2 + k];  // sub-range keys: digits of G
      run.from_lo[k] = 0;
    }
  }
  g_last_plan.store(hybrid ? (t_xf.on ? 3 : 2) : (prefix_shift >= 0 ? 1 : 0), std::memory_order_relaxed);
  const bool uchk = prefix_shift < 0 && !t_xf.on;  // prefix mode was chosen because the digits spread
  switch (kind) {
    case KIND_I64: e = launch_encode<KIND_I64>(has_lo, uchk, col, width, n, desc, hi[0], lo[0], run, ghist, s, 0, 0, too_long); break;
    case KIND_U64: e = launch_encode<KIND_U64>(has_lo, uchk, col, width, n, desc, hi[0], lo[0], run, ghist, s, 0, 0, too_long); break;
    case KIND_F64: e = launch_encode<KIND_F64>(has_lo, uchk, col, width, n, desc, hi[0], lo[0], run, ghist, s, 0, 0, too_long); break;
    case KIND_PHILOX: e = launch_encode<KIND_PHILOX>(false, false, nullptr, 8, n, false, hi[0], lo[0], run, ghist, s, seed, subseq, too_long); break;
    default: e = launch_encode<KIND_FSB>(has_lo, uchk, col, width, n, desc, hi[0], lo[0], run, ghist, s, 0, 0, too_long); break;
  }
  DS_CUDA(e);
  ds::note_launch(); scan_hist_kernel<<<run.count, RADIX, 0, s>>>(ghist, n, bin_base, skip, coll);
  DS_LAUNCH_CHECK();

  uint32_t skip_h[MAX_PASSES] = {0};
  if (prefix_shift < 0 && !t_xf.on && !(options & DS_SORT_NO_PASS_SKIP)) {
    DS_CUDA(cudaMemcpyAsync(skip_h, skip, sizeof(skip_h), cudaMemcpyDeviceToHost, s));
    DS_CUDA(cudaStreamSynchronize(s));  // host sync: 40 bytes (constant digits are skipped)
  }
  int exec[MAX_PASSES], ne = 0;
  for (int p = 0; p < run.count; ++p)
    if (!skip_h[p]) exec[ne++] = p;

  int cur = 0;
  if (ne == 0) {
    ds::note_launch(); iota_kernel<<<(n + 255) / 256, 256, 0, s>>>(idx_out, n);
    DS_LAUNCH_CHECK();
  }
  const uint32_t* idx_src = nullptr;
  for (int j = 0; j < ne; ++j) {
    const int p = exec[j];
    uint32_t* idx_dst = ((ne - 1 - j) % 2 == 0) != hybrid ? idx_out : idx_alt;  // hybrid: last pass -> idx_alt
    DS_CUDA(cudaMemsetAsync(lookback, 0, sizeof(uint32_t) * static_cast<size_t>(tiles) * RADIX, s));
    DS_CUDA(launch_pass(cfg, has_lo, run.from_lo[p] != 0, j == 0, hi[cur], lo[cur], idx_src,
                        hi[cur ^ 1], lo[cur ^ 1], idx_dst, n, run.shift[p], bin_base + p * RADIX,
                        lookback, tile_counters + p, s));
    idx_src = idx_dst;
    cur ^= 1;
  }
  if (hybrid) {
    uint32_t* starts = bucket_starts;
    ds::note_launch();
    bucket_bounds_kernel<<<(n + 1023) / 1024, 256, 0, s>>>(hi[cur], n, starts, t_xf);
    DS_LAUNCH_CHECK();
    int out_kind = LS_OUT_NONE;  // float64 keys: the caller's gather below produces the column
    if (sorted_out != nullptr) {
      if (kind == KIND_FSB) out_kind = LS_OUT_FSB10;
      if (kind == KIND_I64) out_kind = LS_OUT_I64;
      if (kind == KIND_U64) out_kind = LS_OUT_U64;
    }
    static DoneMask attr_done[8];
#define DS_BUCKET_LAUNCH_X(HL, OUT, XF)                                                                \
  do {                                                                                                 \
    DS_CUDA(set_smem_attr(bucket_local_sort_kernel<HL, OUT, XF>, LS_SMEM_BYTES, &attr_done[OUT + (XF ? 4 : 0)])); \
    ds::note_launch();                                                                                 \
    bucket_local_sort_kernel<HL, OUT, XF><<<LS_BUCKETS, LS_THREADS, LS_SMEM_BYTES, s>>>(               \
        hi[cur], lo[cur], idx_src, starts, desc, static_cast<uint8_t*>(sorted_out), idx_out, too_long, t_xf); \
  } while (0)
#define DS_BUCKET_LAUNCH(HL, OUT)          \
  do {                                     \
    if (t_xf.on)                           \
      DS_BUCKET_LAUNCH_X(HL, OUT, true);   \
    else                                   \
      DS_BUCKET_LAUNCH_X(HL, OUT, false);  \
  } while (0)
    switch (out_kind) {
      case LS_OUT_FSB10: DS_BUCKET_LAUNCH(true, LS_OUT_FSB10); break;
      case LS_OUT_I64: DS_BUCKET_LAUNCH(false, LS_OUT_I64); break;
      case LS_OUT_U64: DS_BUCKET_LAUNCH(false, LS_OUT_U64); break;
      default: DS_BUCKET_LAUNCH(false, LS_OUT_NONE); break;
    }
#undef DS_BUCKET_LAUNCH
#undef DS_BUCKET_LAUNCH_X
    DS_LAUNCH_CHECK();
    uint32_t flag = 0;
    DS_CUDA(cudaMemcpyAsync(&flag, too_long, sizeof(flag), cudaMemcpyDeviceToHost, s));
    DS_CUDA(cudaStreamSynchronize(s));  // host sync #2: 4 bytes
    if (flag)  // a bucket or a run did not fit: the all-global prefix plan handles any data
      return sort_indices_impl(kind, col, width, n64, descending, idx_out, sorted_out,
                               options | DS_SORT_NO_HYBRID, ws, ws_bytes, stream, fn, seed, subseq);
    if (kind == KIND_F64 && sorted_out != nullptr) return gather_rows_impl(col, 8, n, idx_out, n, sorted_out, s);
    return DS_OK;
  }
  const bool fused_decode = prefix_shift >= 0 && kind == KIND_FSB && width == 10 && sorted_out != nullptr &&
                            (reinterpret_cast<uintptr_t>(sorted_out) & 15) == 0 && !g_no_fused_decode.load(std::memory_order_relaxed);
  if (fused_decode) {
    // the other half of the ping-pong buffers is dead now: it holds the list of run heads
    uint32_t* head_list = reinterpret_cast<uint32_t*>(hi[cur ^ 1]);
    uint32_t* head_count = too_long + 1;  // zeroed with the rest of ctrl
    ds::note_launch();
    decode10_find_runs_kernel<<<(n + DECODE_ROWS - 1) / DECODE_ROWS, 256, 0, s>>>(
        hi[cur], lo[cur], n, desc, prefix_shift, static_cast<uint8_t*>(sorted_out), head_list, head_count);
    DS_LAUNCH_CHECK();
    ds::note_launch();
    repair_listed_runs_kernel<<<sm_count() * 4, 256, 0, s>>>(hi[cur], lo[cur], idx_out, n, prefix_shift, head_list,
                                                             head_count, desc, static_cast<uint8_t*>(sorted_out),
                                                             too_long);
    DS_LAUNCH_CHECK();
    uint32_t flag = 0;
    DS_CUDA(cudaMemcpyAsync(&flag, too_long, sizeof(flag), cudaMemcpyDeviceToHost, s));
    DS_CUDA(cudaStreamSynchronize(s));  // host sync #2 (prefix mode only): 4 bytes
    if (flag)
      return sort_indices_impl(kind, col, width, n64, descending, idx_out, sorted_out,
                               options | DS_SORT_NO_PREFIX_REPAIR, ws, ws_bytes, stream, fn, seed, subseq);
    return DS_OK;
  }
  if (prefix_shift >= 0) {
    ds::note_launch();
    const uint32_t repair_blocks = (n + 256 * REPAIR_VEC - 1) / (256 * REPAIR_VEC);
    if (has_lo)
      repair_runs_kernel<true><<<repair_blocks, 256, 0, s>>>(hi[cur], lo[cur], idx_out, n, prefix_shift, too_long);
    else
      repair_runs_kernel<false><<<repair_blocks, 256, 0, s>>>(hi[cur], lo[cur], idx_out, n, prefix_shift, too_long);
    DS_LAUNCH_CHECK();
    uint32_t flag = 0;
    DS_CUDA(cudaMemcpyAsync(&flag, too_long, sizeof(flag), cudaMemcpyDeviceToHost, s));
    DS_CUDA(cudaStreamSynchronize(s));  // host sync #2 (prefix mode only): 4 bytes
    if (flag)  // a long run of equal prefixes: the estimate was wrong, sort all digits
      return sort_indices_impl(kind, col, width, n64, descending, idx_out, sorted_out,
                               options | DS_SORT_NO_PREFIX_REPAIR, ws, ws_bytes, stream, fn, seed, subseq);
  }
  if (sorted_out != nullptr) {
    const int threads = 256, blocks = (n + threads - 1) / threads;
    switch (kind) {
      case KIND_I64: ds::note_launch(); decode_kernel<KIND_I64><<<blocks, threads, 0, s>>>(hi[cur], lo[cur], width, n, desc, sorted_out); break;
      case KIND_U64: ds::note_launch(); decode_kernel<KIND_U64><<<blocks, threads, 0, s>>>(hi[cur], lo[cur], width, n, desc, sorted_out); break;
      case KIND_F64: return gather_rows_impl(col, 8, n, idx_out, n, sorted_out, s);
      default:
        ds::note_launch();
        if ((reinterpret_cast<uintptr_t>(sorted_out) & 15) == 0) {
          const int sblocks = static_cast<int>((n + DECODE_ROWS - 1) / DECODE_ROWS);
          if (width == 10)
            decode_fsb_staged_kernel<true><<<sblocks, 256, 0, s>>>(hi[cur], lo[cur], width, n, desc,
                                                                   static_cast<uint8_t*>(sorted_out));
          else
            decode_fsb_staged_kernel<false><<<sblocks, 256, 0, s>>>(hi[cur], lo[cur], width, n, desc,
                                                                    static_cast<uint8_t*>(sorted_out));
        } else
          decode_kernel<KIND_FSB><<<blocks, threads, 0, s>>>(hi[cur], lo[cur], width, n, desc, sorted_out);
        break;
    }
    DS_LAUNCH_CHECK();
  }
  return DS_OK;
}

}  // namespace ds

extern "C" {

ds_status ds_sort_indices_i64(const int64_t* col, uint64_t n, int32_t descending, uint32_t* idx_out,
                              int64_t* sorted_out, int32_t options, void* ws, size_t* ws_bytes,
                              void* stream) {
  return ds::sort_indices_impl(ds::KIND_I64, col, 8, n, descending, idx_out, sorted_out, options, ws,
                               ws_bytes, stream, __func__, 0, 0);
}

ds_status ds_sort_indices_u64(const uint64_t* col, uint64_t n, int32_t descending, uint32_t* idx_out,
                              uint64_t* sorted_out, int32_t options, void* ws, size_t* ws_bytes,
                              void* stream) {
  return ds::sort_indices_impl(ds::KIND_U64, col, 8, n, descending, idx_out, sorted_out, options, ws,
                               ws_bytes, stream, __func__, 0, 0);
}

ds_status ds_sort_indices_f64(const double* col, uint64_t n, int32_t descending, uint32_t* idx_out,
                              double* sorted_out, int32_t options, void* ws, size_t* ws_bytes,
                              void* stream) {
  return ds::sort_indices_impl(ds::KIND_F64, col, 8, n, descending, idx_out, sorted_out, options, ws,
                               ws_bytes, stream, __func__, 0, 0);
}

ds_status ds_sort_indices_fsb(const uint8_t* col, uint32_t width, uint64_t n, int32_t descending,
                              uint32_t* idx_out, uint8_t* sorted_out, int32_t options, void* ws,
                              size_t* ws_bytes, void* stream) {
  return ds::sort_indices_impl(ds::KIND_FSB, col, width, n, descending, idx_out, sorted_out, options,
                               ws, ws_bytes, stream, __func__, 0, 0);
}

ds_status ds_random_permutation(uint64_t seed, uint64_t subsequence, uint64_t n, uint32_t* perm_out,
                                void* ws, size_t* ws_bytes, void* stream) {
  // Philox keys are produced inside the encode kernel (nothing is read from HBM);
  // colliding 64-bit keys are ordered by row index, which keeps the result a permutation.
  return ds::sort_indices_impl(ds::KIND_PHILOX, nullptr, 8, n, 0, perm_out, nullptr, 0, ws,
                               ws_bytes, stream, __func__, seed, subsequence);
}

// Tuning hook used by bench.py --sweep; not part of the reference-facing surface.
// 0: prefix mode runs repair_runs_kernel + the decode kernel separately (the older path);
// 1 (default): the fused decode + run-head list pair.  Returns the previous setting.
int32_t ds_debug_set_fused_decode(int32_t enable) {
  const int old = ds::g_no_fused_decode.load() ? 0 : 1;
  ds::g_no_fused_decode.store(enable == 0);
  return old;
}

// Hybrid plan for 10-byte keys (two global passes + one CTA per 16-bit bucket): 0 never,
// 1 (default) when n / 65536 is between 400 and 1650 rows, 2 also for smaller n (tests).
// Returns the previous mode.
int32_t ds_debug_set_hybrid(int32_t mode) {
  const int old = ds::g_hybrid_mode.load();
  if (mode >= 0 && mode <= 2) ds::g_hybrid_mode.store(mode);
  return old;
}

// The hybrid plan on keys that fill only a sub-range of the key space (KeyXform): 1 (default)
// on, 0 off.  Returns the previous setting.
int32_t ds_debug_set_subrange(int32_t enable) {
  const int old = ds::g_no_subrange.load() ? 0 : 1;
  ds::g_no_subrange.store(enable == 0);
  return old;
}

// Plan the last sort call of this process ended with (after any fall-back): 0 all passes,
// 1 prefix passes + run repair, 2 hybrid, 3 hybrid on a sub-range of the key space; -1 none yet.
int32_t ds_debug_last_sort_plan(void) { return ds::g_last_plan.load(); }

// Host arithmetic of the sub-range transform, for tests without a GPU: out = {base, span, R, s};
// returns 1 when the transform would be used for n keys whose sample spans [min, max].
int32_t ds_debug_key_xform(uint64_t min, uint64_t max, uint64_t n, uint64_t* out) {
  const ds::KeyXform xf = ds::make_key_xform(min, max, n);
  if (out != nullptr) {
    out[0] = xf.base;
    out[1] = xf.span;
    out[2] = xf.R;
    out[3] = xf.s;
  }
  return static_cast<int32_t>(xf.on);
}

int32_t ds_debug_set_radix_config(int32_t cfg) {
  int old = ds::g_cfg.load();
  if (cfg >= 0 && cfg < ds::NUM_CFGS && ds::cfg_known(cfg)) ds::g_cfg.store(cfg);
  return old;
}

}  // extern "C"
