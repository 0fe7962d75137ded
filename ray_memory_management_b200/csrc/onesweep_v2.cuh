// onesweep pass, second generation: "early counts" + ballot matching.
//
// Differences from onesweep_pass_kernel (radix_sort.cu):
//  * per-warp digit histograms are built first with independent shared-memory atomics
//    (no serial chain), so the tile's digit counts are known -- and PUBLISHED to the
//    look-back array -- before any ranking work starts; successors never wait on our
//    ranking, and our own look-back wait overlaps our ranking + exchange.
//  * peers of a digit are found with 8 vote.ballot instructions (one per digit bit)
//    instead of match.any, which is a long-latency MIO instruction on sm_100a.
//  * the running per-warp offset is advanced by the peer-group leader with one
//    shared-memory atomicAdd and broadcast with a shuffle; there is no __syncwarp()
//    pair per item, so the 8..16 items of a thread pipeline.
// Stability argument is unchanged: a warp owns a contiguous span of the tile in
// warp-striped order, items are visited in position order, peers ranked by lane.
#pragma once

namespace ds {

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

// FULL: every tile of this launch is complete (the ragged last tile is a second,
// one-block launch with FULL=false that continues the same tile counter), which
// removes all per-item bounds predicates.  USE_MATCH: match.any instead of 8 ballots.
template <int THREADS, int ITEMS, int MIN_BLOCKS, bool HAS_LO, bool DIGIT_LO, bool GEN_IDX,
          bool FULL, bool USE_MATCH>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
    onesweep_pass_kernel_v2(const uint64_t* __restrict__ hi_in, const uint16_t* __restrict__ lo_in,
                            const uint32_t* __restrict__ idx_in, uint64_t* __restrict__ hi_out,
                            uint16_t* __restrict__ lo_out, uint32_t* __restrict__ idx_out,
                            uint32_t n, int shift, const uint32_t* __restrict__ bin_base,
                            volatile uint32_t* lookback, uint32_t* tile_counter) {
  static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit is required");
  using S = PassSmem<THREADS, ITEMS, HAS_LO>;
  constexpr int TILE = S::TILE;
  constexpr int WARPS = S::WARPS;
  extern __shared__ __align__(16) unsigned char smem[];
  uint32_t* warp_hist = reinterpret_cast<uint32_t*>(smem);  // [WARPS][RADIX], then reused:
  uint64_t* ex_hi = reinterpret_cast<uint64_t*>(smem);      // [TILE]
  uint32_t* ex_idx = reinterpret_cast<uint32_t*>(smem + TILE * 8);
  uint16_t* ex_lo = reinterpret_cast<uint16_t*>(smem + TILE * 12);
  __shared__ uint32_t s_gofs[RADIX];
  __shared__ uint32_t s_dstart[RADIX];
  __shared__ uint32_t s_warp_tot[8];
  __shared__ uint32_t s_tile;

  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const uint32_t lane = tid & 31;
  if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
#pragma unroll
  for (int i = tid; i < WARPS * RADIX; i += THREADS) warp_hist[i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint32_t tile_base = tile * TILE;
  const uint32_t tile_valid = min(static_cast<uint32_t>(TILE), n - tile_base);
  const bool full = FULL || tile_valid == TILE;

  uint64_t hi[ITEMS];
  uint32_t lo[ITEMS];
  uint32_t idx[ITEMS];
  uint32_t pos[ITEMS];
  const uint32_t wofs = warp * 32 * ITEMS + lane;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const uint32_t lp = wofs + i * 32;
    const uint32_t gp = tile_base + lp;
    if (full || lp < tile_valid) {
      hi[i] = ld_stream(hi_in + gp);
      lo[i] = HAS_LO ? ld_stream(lo_in + gp) : 0;
      idx[i] = GEN_IDX ? gp : ld_stream(idx_in + gp);
    } else {
      hi[i] = 0;
      lo[i] = 0;
      idx[i] = 0;
    }
  }

  // ---- A. early counts: independent atomics into this warp's histogram
  uint32_t* wh = warp_hist + warp * RADIX;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    if (full || (wofs + i * 32 < tile_valid))
      atomicAdd(&wh[digit_of<HAS_LO, DIGIT_LO>(hi[i], lo[i], shift)], 1u);
  }
  __syncthreads();

  // ---- B. per digit: tile count -> publish the aggregate at once; scan over digits; then
  // turn each warp's count into its running offset INCLUDING the digit's tile start, so
  // that ranking needs no second table lookup.
  volatile uint32_t* my = lookback + static_cast<size_t>(tile) * RADIX + tid;
  uint32_t my_sum = 0;
  if (tid < RADIX) {
    uint32_t sum = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) sum += warp_hist[w * RADIX + tid];
    st_relaxed_gpu(my, (tile == 0 ? LB_INCLUSIVE : LB_AGGREGATE) | sum);
    my_sum = sum;
    uint32_t v = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
      if ((int)lane >= o) v += t;
    }
    if (lane == 31) s_warp_tot[warp] = v;
    asm volatile("bar.sync 1, 256;");
    uint32_t add = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w)
      if (w < warp) add += s_warp_tot[w];
    uint32_t run = v + add - sum;  // first tile-local position of this digit
    s_dstart[tid] = run;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = warp_hist[w * RADIX + tid];
      warp_hist[w * RADIX + tid] = run;
      run += c;
    }
  }
  __syncthreads();

  // ---- C. rank: ballots find the peers, the leader advances the warp's offset
  const uint32_t lt_mask = lanemask_lt();
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const bool valid = full || (wofs + i * 32 < tile_valid);
    const uint32_t d = digit_of<HAS_LO, DIGIT_LO>(hi[i], lo[i], shift);
    uint32_t peers;
    if (USE_MATCH) {
      peers = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
    } else {
      peers = ballot_match8(d);
      if (!FULL && !full) peers &= __ballot_sync(0xffffffffu, valid);
    }
    const uint32_t lt = __popc(peers & lt_mask);
    uint32_t base = 0;
    if (valid && lt == 0) base = atomicAdd(&wh[d], __popc(peers));
    base = __shfl_sync(0xffffffffu, base, __ffs(peers) - 1);
    pos[i] = base + lt;
  }
  __syncthreads();  // warp_hist is dead from here: the exchange buffers overlay it

#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    if (full || (wofs + i * 32 < tile_valid)) {
      ex_hi[pos[i]] = hi[i];
      ex_idx[pos[i]] = idx[i];
      if (HAS_LO) ex_lo[pos[i]] = static_cast<uint16_t>(lo[i]);
    }
  }

  // ---- D. look back over earlier tiles (their aggregates were published in their B)
  if (tid < RADIX) {
    uint32_t excl = 0;
    if (tile != 0) {
      // walk back over predecessor tiles, LB_BATCH independent loads in flight at a time
      constexpr int LB_BATCH = 4;
      int t = static_cast<int>(tile) - 1;
      bool done = false;
      while (!done) {
        uint32_t s[LB_BATCH];
#pragma unroll
        for (int k = 0; k < LB_BATCH; ++k)
          s[k] = (t - k >= 0) ? ld_relaxed_gpu(lookback + static_cast<size_t>(t - k) * RADIX + tid) : LB_INCLUSIVE;
#pragma unroll
        for (int k = 0; k < LB_BATCH; ++k) {
          if (done) break;
          if ((s[k] & ~LB_VALUE) == 0) break;  // not published yet: re-read from here
          excl += s[k] & LB_VALUE;
          --t;
          if (s[k] & LB_INCLUSIVE) done = true;
        }
      }
      st_relaxed_gpu(my, LB_INCLUSIVE | (excl + my_sum));
    }
    s_gofs[tid] = bin_base[tid] + excl - s_dstart[tid];
  }
  __syncthreads();

  // ---- E. coalesced scatter
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const uint32_t p = tid + i * THREADS;
    if (full || p < tile_valid) {
      const uint64_t h = ex_hi[p];
      const uint32_t l = HAS_LO ? ex_lo[p] : 0;
      const uint32_t x = ex_idx[p];
      const uint32_t d = digit_of<HAS_LO, DIGIT_LO>(h, l, shift);
      const uint32_t dst = s_gofs[d] + p;
      hi_out[dst] = h;
      if (HAS_LO) lo_out[dst] = static_cast<uint16_t>(l);
      idx_out[dst] = x;
    }
  }
}

}  // namespace ds
