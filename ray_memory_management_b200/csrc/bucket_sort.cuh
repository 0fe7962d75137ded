// Hybrid plan (10-byte keys and 64-bit keys): two global onesweep passes on the two leading bytes leave the
// rows grouped into 65536 buckets (about n / 65536 rows each, contiguous, input order kept
// inside a bucket).  One CTA per bucket then finishes the job without further global passes:
//   * one counting sort on key bits [36, 48) (4096 bins for ~1500 rows) entirely in shared
//     memory, positions handed out by shared-memory atomics,
//   * ordering of the short runs of rows that share a bin by (full key, position on entry),
//   * decode to the 10-byte key column and a coalesced store of keys and row ids.
// This replaces two global passes, the repair kernel and the decode kernel of the plain
// prefix plan.  A bucket that does not fit the CTA (skewed leading bytes) or a run longer than
// the repair limit raises *too_long and the caller falls back to the all-global plan.
#pragma once

namespace ds {

constexpr int LS_THREADS = 256;
constexpr int LS_ITEMS = 8;
constexpr int LS_CAP = LS_THREADS * LS_ITEMS;  // rows a CTA can hold
constexpr int LS_WARPS = LS_THREADS / 32;
constexpr uint32_t LS_BUCKETS = 65536;
constexpr int LS_BINS = 4096;                  // bins of the CTA's counting sort ...
constexpr int LS_BIN_SHIFT = 36;               // ... = key bits [36, 48)
constexpr int LS_SMEM_BYTES = LS_CAP * 16 + LS_CAP * 10 + 16;

// starts[b] = first row whose leading 16 bits are >= b, starts[65536] = n (empty buckets get the
// next bucket's start).  `hi` is sorted by those bits.
__global__ void __launch_bounds__(256)
    bucket_bounds_kernel(const uint64_t* __restrict__ hi, uint32_t n, uint32_t* __restrict__ starts,
                         const KeyXform xf) {
  const uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i0 >= n) return;
  uint32_t b[5];  // b[k + 1] = bucket of row i0 + k, b[0] = bucket of row i0 - 1
  if (i0 + 4 <= n) {
    const ulonglong2 x = *reinterpret_cast<const ulonglong2*>(hi + i0);
    const ulonglong2 y = *reinterpret_cast<const ulonglong2*>(hi + i0 + 2);
    b[1] = g28(x.x, xf) >> 12;
    b[2] = g28(x.y, xf) >> 12;
    b[3] = g28(y.x, xf) >> 12;
    b[4] = g28(y.y, xf) >> 12;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) b[k + 1] = i0 + k < n ? g28(hi[i0 + k], xf) >> 12 : 0;
  }
  if (i0 == 0) {
    for (uint32_t v = 0; v <= b[1]; ++v) starts[v] = 0;
    b[0] = b[1];
  } else {
    b[0] = g28(hi[i0 - 1], xf) >> 12;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t i = i0 + k;
    if (i >= n) break;
    for (uint32_t v = b[k] + 1; v <= b[k + 1]; ++v) starts[v] = i;  // usually zero or one iteration
    if (i == n - 1)
      for (uint32_t v = b[k + 1] + 1; v <= LS_BUCKETS; ++v) starts[v] = n;
  }
}

// OUT: what is written besides the row ids -- LS_OUT_FSB10: the 10-byte key column (HAS_LO);
// LS_OUT_I64 / LS_OUT_U64: the 64-bit key column decoded from `hi`; LS_OUT_NONE: nothing
// (permutations; float64 keys, whose column is gathered by the caller).
enum { LS_OUT_NONE = 0, LS_OUT_FSB10 = 1, LS_OUT_I64 = 2, LS_OUT_U64 = 3 };

template <bool HAS_LO, int OUT, bool XF>
__global__ void __launch_bounds__(LS_THREADS, 4)
    bucket_local_sort_kernel(const uint64_t* __restrict__ hi_in, const uint16_t* __restrict__ lo_in,
                             const uint32_t* __restrict__ idx_in, const uint32_t* __restrict__ starts,
                             bool desc, uint8_t* __restrict__ keys_out, uint32_t* __restrict__ idx_out,
                             uint32_t* __restrict__ too_long, const KeyXform xf) {
  static_assert(OUT != LS_OUT_FSB10 || HAS_LO, "10-byte keys carry their last two bytes in lo");
  // The CTA's ~1500 rows are spread over LS_BINS bins by key bits [LS_BIN_SHIFT, 48) with one
  // counting sort whose scatter takes positions from shared-memory atomics -- NOT stable, so
  // every row also carries its position in the bucket (the stable order so far), and the runs
  // of rows that share a bin (about a third of the rows, runs of 2-4) are then ordered by
  // (full key, that position).  Rows of different bins never tie on the leading 32 bits' worth
  // of order that the bins already decide, so the result is the stable full-key order.
  // Shared-memory wavefronts are what bounds this kernel (ncu: l1tex 91 % busy, half of the
  // wavefronts bank conflicts), so a row is two 8-byte words -- the key's hi word and a meta
  // word (lo:16 | entry position:16 | row id:32, compared as one integer it orders by (lo,
  // position)) -- which halves the scattered stores and the moves of the insertion sort.
  extern __shared__ __align__(16) unsigned char ls_dyn[];  // LS_SMEM_BYTES (over the 48 KB static limit)
  uint64_t* ex_hi = reinterpret_cast<uint64_t*>(ls_dyn);                  // rows by position
  uint64_t* ex_meta = reinterpret_cast<uint64_t*>(ls_dyn + LS_CAP * 8);
  uint8_t* stage = ls_dyn + LS_CAP * 16;             // key bytes as they lie in the output ...
  uint32_t* hist = reinterpret_cast<uint32_t*>(stage);  // ... and, before that, the bin counters
  __shared__ uint32_t s_warp_tot[LS_WARPS];
  static_assert(LS_BINS * 4 <= LS_CAP * 10 + 16, "the bin counters overlay the staging buffer");

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const uint32_t lane = tid & 31;
  const uint32_t s = starts[blockIdx.x];
  const uint32_t R = starts[blockIdx.x + 1] - s;
  if (R == 0) return;
  if (R > LS_CAP) {
    if (tid == 0) atomicOr(too_long, 1u);
    return;
  }

  uint64_t hi[LS_ITEMS], meta[LS_ITEMS];
  const uint32_t wofs = warp * 32 * LS_ITEMS + lane;
#pragma unroll
  for (int i = 0; i < LS_ITEMS; ++i) {
    const uint32_t lp = wofs + i * 32;
    if (lp < R) {
      hi[i] = ld_stream(hi_in + s + lp);
      const uint64_t l = HAS_LO ? ld_stream(lo_in + s + lp) : 0;
      meta[i] = (l << 48) | (static_cast<uint64_t>(lp) << 32) | ld_stream(idx_in + s + lp);
    } else {
      hi[i] = 0;
      meta[i] = 0;
    }
  }
  for (int i = tid; i < LS_BINS; i += LS_THREADS) hist[i] = 0;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < LS_ITEMS; ++i)
    if (wofs + i * 32 < R) atomicAdd(&hist[g28t<XF>(hi[i], xf) & (LS_BINS - 1)], 1u);
  __syncthreads();
  {  // exclusive scan of the bin counts: LS_BINS / LS_THREADS consecutive bins per thread
    constexpr int PER = LS_BINS / LS_THREADS;
    uint32_t c[PER];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < PER; k += 4) {
      const uint4 q = *reinterpret_cast<const uint4*>(hist + tid * PER + k);
      c[k] = q.x; c[k + 1] = q.y; c[k + 2] = q.z; c[k + 3] = q.w;
      sum += q.x + q.y + q.z + q.w;
    }
    uint32_t v = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
      if ((int)lane >= o) v += t;
    }
    if (lane == 31) s_warp_tot[warp] = v;
    __syncthreads();
    uint32_t run = v - sum;
    for (int w = 0; w < warp; ++w) run += s_warp_tot[w];
#pragma unroll
    for (int k = 0; k < PER; k += 4) {
      uint4 q;
      q.x = run; run += c[k];
      q.y = run; run += c[k + 1];
      q.z = run; run += c[k + 2];
      q.w = run; run += c[k + 3];
      *reinterpret_cast<uint4*>(hist + tid * PER + k) = q;
    }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < LS_ITEMS; ++i) {
    if (wofs + i * 32 < R) {
      const uint32_t p = atomicAdd(&hist[g28t<XF>(hi[i], xf) & (LS_BINS - 1)], 1u);
      ex_hi[p] = hi[i];
      ex_meta[p] = meta[i];
    }
  }
  __syncthreads();

  // ---- order the runs of rows that share a bin by (key, position on entry): the thread that
  // holds a run's first row sorts it by insertion.  (Letting every member count its rank and
  // move after a barrier was measured slower: 1.41 against 1.20 ms.)
  // There is no barrier between finding the heads and sorting: a thread may read a neighbouring
  // run's ex_hi words (head test, scan for the run's end) while that run's owner moves them.
  // What it compares is G only, which is equal for all members of a run by definition, an
  // insertion sort only permutes words inside its run, and aligned 64-bit shared-memory stores
  // do not tear -- so every value it can see gives the same answer.  Runs never overlap and only
  // the owner writes.  (A racecheck tool reports these reads; they are intended.)
  {
    uint32_t heads = 0;
#pragma unroll
    for (int i = 0; i < LS_ITEMS; ++i) {  // one load per row; the neighbours' bins come by shuffle
      const uint32_t p = tid + i * LS_THREADS;
      const uint32_t none = 0xffffffffu;   // bins have 28 bits
      const uint32_t mine = p < R ? g28t<XF>(ex_hi[p], xf) : none;
      uint32_t next = __shfl_down_sync(0xffffffffu, mine, 1);
      uint32_t prev = __shfl_up_sync(0xffffffffu, mine, 1);
      if (lane == 31) next = p + 1 < R ? g28t<XF>(ex_hi[p + 1], xf) : none;
      if (lane == 0) prev = (p > 0 && p < R) ? g28t<XF>(ex_hi[p - 1], xf) : none;
      if (p + 1 < R && next == mine && prev != mine) heads |= 1u << i;
    }
    while (heads) {
      const int i = __ffs(heads) - 1;
      heads &= heads - 1;
      const uint32_t p = tid + i * LS_THREADS;
      const uint32_t bin = g28t<XF>(ex_hi[p], xf);
      uint32_t len = 2;
      while (p + len < R && len <= REPAIR_MAX_RUN && g28t<XF>(ex_hi[p + len], xf) == bin) ++len;
      if (len > REPAIR_MAX_RUN) {
        atomicOr(too_long, 1u);
        continue;
      }
      for (uint32_t a = 1; a < len; ++a) {  // insertion sort inside shared memory
        const uint64_t kh = ex_hi[p + a];
        const uint64_t km = ex_meta[p + a];
        uint32_t b = a;
        while (b > 0) {
          const uint64_t qh = ex_hi[p + b - 1];
          if (qh < kh) break;
          const uint64_t qm = ex_meta[p + b - 1];
          if (qh == kh && (qm >> 32) < (km >> 32)) break;
          ex_hi[p + b] = qh;
          ex_meta[p + b] = qm;
          --b;
        }
        ex_hi[p + b] = kh;
        ex_meta[p + b] = km;
      }
    }
  }
  __syncthreads();  // also: the bin counters are dead, the staging buffer may overwrite them

  if (OUT != LS_OUT_FSB10) {  // 64-bit keys (or none): plain coalesced stores
#pragma unroll
    for (int i = 0; i < LS_ITEMS; ++i) {
      const uint32_t p = tid + i * LS_THREADS;
      if (p < R) {
        idx_out[s + p] = static_cast<uint32_t>(ex_meta[p]);
        const uint64_t h0 = XF ? (ex_hi[p] >> xf.s) + xf.base : ex_hi[p];  // the key's own encoding
        if (OUT == LS_OUT_I64) reinterpret_cast<int64_t*>(keys_out)[s + p] = dec_i64(h0, desc);
        if (OUT == LS_OUT_U64) reinterpret_cast<uint64_t*>(keys_out)[s + p] = desc ? ~h0 : h0;
      }
    }
    return;
  }
  // ---- decode into the staging buffer laid out like the output, then coalesced stores.
  // A row's 10 bytes start on an even address: two 4-byte stores and one 2-byte store, in
  // the order its alignment allows.
  const uint64_t out_byte = static_cast<uint64_t>(s) * 10u;
  const uint32_t mis = static_cast<uint32_t>(out_byte & 15u);  // even: rows are 10 bytes
#pragma unroll
  for (int i = 0; i < LS_ITEMS; ++i) {
    const uint32_t p = tid + i * LS_THREADS;
    if (p < R) {
      const uint64_t m = ex_meta[p];
      const uint64_t h0 = XF ? (ex_hi[p] >> xf.s) + xf.base : ex_hi[p];  // the key's own encoding
      const uint64_t hh = desc ? ~h0 : h0;
      const uint32_t ll = static_cast<uint32_t>((desc ? ~m : m) >> 48);
      const uint64_t be = bswap64(hh);                                   // bytes 0..7 of the row
      const uint32_t t2 = ((ll >> 8) | ((ll & 0xffu) << 8)) & 0xffffu;   // bytes 8..9
      const uint32_t o = mis + p * 10u;
      if ((o & 3u) == 0) {
        *reinterpret_cast<uint32_t*>(stage + o) = static_cast<uint32_t>(be);
        *reinterpret_cast<uint32_t*>(stage + o + 4) = static_cast<uint32_t>(be >> 32);
        *reinterpret_cast<uint16_t*>(stage + o + 8) = static_cast<uint16_t>(t2);
      } else {
        *reinterpret_cast<uint16_t*>(stage + o) = static_cast<uint16_t>(be);
        *reinterpret_cast<uint32_t*>(stage + o + 2) = static_cast<uint32_t>(be >> 16);
        *reinterpret_cast<uint32_t*>(stage + o + 6) = static_cast<uint32_t>(be >> 48) | (t2 << 16);
      }
      idx_out[s + p] = static_cast<uint32_t>(m);
    }
  }
  __syncthreads();
  uint8_t* g = keys_out + (out_byte - mis);  // 16-byte aligned
  const uint32_t end = mis + R * 10u;
  const uint32_t first_vec = (mis + 15u) >> 4, last_vec = end >> 4;
  for (uint32_t v = first_vec + tid; v < last_vec; v += LS_THREADS)
    st_stream(reinterpret_cast<uint4*>(g) + v, reinterpret_cast<const uint4*>(stage)[v]);
  const uint32_t head_end = min(first_vec << 4, end);
  for (uint32_t o = mis + tid; o < head_end; o += LS_THREADS) g[o] = stage[o];
  const uint32_t tail_begin = max(last_vec << 4, head_end);
  for (uint32_t o = tail_begin + tid; o < end; o += LS_THREADS) g[o] = stage[o];
}

}  // namespace ds
