// onesweep pass, third generation (the first two were retired after round 1; their profiles are
// in profiles/r01_pass_kernel_v2_summary.txt).  Early per-warp digit counts, tile aggregate
// published before ranking, batched look-back, full-tile fast path; compared with the second
// generation the two per-item costs that dominated its profile are replaced:
//
//  * peer matching through shared memory instead of 8 ballots.  On B200 the ALU pipe was
//    the busiest unit of v2 (~70 ALU instructions per key, most of them the ballot / LOP3
//    chains).  Here every lane ORs its lane bit into a per-warp 256-entry table at its
//    digit, reads the entry back (= the mask of lanes holding the same digit) and clears
//    it: 3 shared-memory instructions instead of ~40 ALU ones.  Two tables alternate
//    between consecutive items so that one __syncwarp() pair per item is enough.
//  * (key, low key, row index) travel through the exchange buffer as ONE 16-byte slot:
//    one STS.128 and one LDS.128 per key instead of three narrower accesses each way.
//
// TABLE / PACKED are template switches so that bench.py can time every combination.
#pragma once

namespace ds {

template <int THREADS, int ITEMS, bool PACKED, bool HAS_LO>
struct PassSmemV3 {
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr int WARPS = THREADS / 32;
  static constexpr int EXCH = TILE * (PACKED ? 16 : (8 + 4 + (HAS_LO ? 2 : 0)));
  static constexpr int HIST = WARPS * RADIX * 4 * 3;  // running offsets + two match tables
  static constexpr int BYTES = (EXCH > HIST ? EXCH : HIST);
};

template <int THREADS, int ITEMS, int MIN_BLOCKS, bool HAS_LO, bool DIGIT_LO, bool GEN_IDX,
          bool FULL, bool TABLE, bool PACKED, bool XF>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
    onesweep_pass_kernel_v3(const uint64_t* __restrict__ hi_in, const uint16_t* __restrict__ lo_in,
                            const uint32_t* __restrict__ idx_in, uint64_t* __restrict__ hi_out,
                            uint16_t* __restrict__ lo_out, uint32_t* __restrict__ idx_out,
                            uint32_t n, int shift, const uint32_t* __restrict__ bin_base,
                            volatile uint32_t* lookback, uint32_t* tile_counter, const KeyXform xf) {
  static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit is required");
  using S = PassSmemV3<THREADS, ITEMS, PACKED, HAS_LO>;
  constexpr int TILE = S::TILE;
  constexpr int WARPS = S::WARPS;
  extern __shared__ __align__(16) unsigned char smem[];
  uint32_t* warp_hist = reinterpret_cast<uint32_t*>(smem);             // [WARPS][RADIX]
  uint32_t* match_tab = warp_hist + WARPS * RADIX;                      // [WARPS][2][RADIX]
  uint4* ex = reinterpret_cast<uint4*>(smem);                           // PACKED: [TILE]
  uint64_t* ex_hi = reinterpret_cast<uint64_t*>(smem);                  // else SoA
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
  for (int i = tid; i < WARPS * RADIX * (TABLE ? 3 : 1); i += THREADS) warp_hist[i] = 0;
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

  // ---- A. early counts
  uint32_t* wh = warp_hist + warp * RADIX;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    if (full || (wofs + i * 32 < tile_valid))
      atomicAdd(&wh[digit_x<HAS_LO, DIGIT_LO, XF>(hi[i], lo[i], shift, xf)], 1u);
  }
  __syncthreads();

  // ---- B. tile counts -> publish; scan over digits; counts -> running offsets
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
    uint32_t run = v + add - sum;
    s_dstart[tid] = run;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = warp_hist[w * RADIX + tid];
      warp_hist[w * RADIX + tid] = run;
      run += c;
    }
  }
  __syncthreads();

  // ---- C. rank
  const uint32_t lt_mask = lanemask_lt();
  const uint32_t lane_bit = 1u << lane;
  volatile uint32_t* mt = match_tab + warp * 2 * RADIX;
#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    const bool valid = full || (wofs + i * 32 < tile_valid);
    const uint32_t d = digit_x<HAS_LO, DIGIT_LO, XF>(hi[i], lo[i], shift, xf);
    uint32_t peers;
    if (TABLE) {
      volatile uint32_t* slot = mt + (i & 1) * RADIX + d;
      if (valid) atomicOr(const_cast<uint32_t*>(slot), lane_bit);
      __syncwarp();
      peers = valid ? *slot : lane_bit;
      __syncwarp();
      if (valid) *slot = 0;  // ordered before the next use of this table by the syncs of item i+1
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
  __syncthreads();  // histograms and tables are dead: the exchange buffer overlays them

#pragma unroll
  for (int i = 0; i < ITEMS; ++i) {
    if (full || (wofs + i * 32 < tile_valid)) {
      if (PACKED) {
        ex[pos[i]] = make_uint4(static_cast<uint32_t>(hi[i]), static_cast<uint32_t>(hi[i] >> 32),
                                idx[i], lo[i]);
      } else {
        ex_hi[pos[i]] = hi[i];
        ex_idx[pos[i]] = idx[i];
        if (HAS_LO) ex_lo[pos[i]] = static_cast<uint16_t>(lo[i]);
      }
    }
  }

  // ---- D. look back over earlier tiles
  if (tid < RADIX) {
    uint32_t excl = 0;
    if (tile != 0) {
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
          if ((s[k] & ~LB_VALUE) == 0) break;
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
      uint64_t h;
      uint32_t l, x;
      if (PACKED) {
        const uint4 v = ex[p];
        h = (static_cast<uint64_t>(v.y) << 32) | v.x;
        x = v.z;
        l = v.w;
      } else {
        h = ex_hi[p];
        l = HAS_LO ? ex_lo[p] : 0;
        x = ex_idx[p];
      }
      const uint32_t d = digit_x<HAS_LO, DIGIT_LO, XF>(h, l, shift, xf);
      const uint32_t dst = s_gofs[d] + p;
      hi_out[dst] = h;
      if (HAS_LO) lo_out[dst] = static_cast<uint16_t>(l);
      idx_out[dst] = x;
    }
  }
}

}  // namespace ds
