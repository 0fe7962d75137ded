// Key loading / radix-order encoding shared by the sort, search and partition kernels.
#pragma once
#include "ds_common.cuh"

namespace ds {

enum KeyKind { KIND_I64 = 0, KIND_U64 = 1, KIND_F64 = 2, KIND_FSB = 3, KIND_PHILOX = 4 };

// Philox4x32-10 (Salmon et al., SC'11): counter = (row, subsequence), key = seed.
// Returns the first 64 bits of the 128-bit output block.
__device__ __forceinline__ uint64_t philox64(uint64_t seed, uint64_t subsequence, uint64_t row) {
  uint32_t c0 = static_cast<uint32_t>(row), c1 = static_cast<uint32_t>(row >> 32);
  uint32_t c2 = static_cast<uint32_t>(subsequence), c3 = static_cast<uint32_t>(subsequence >> 32);
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    const uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
    c0 = n0; c1 = l1; c2 = n2; c3 = l0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return (static_cast<uint64_t>(c1) << 32) | c0;
}

template <int KIND>
__device__ __forceinline__ void load_key(const void* col, uint32_t width, uint64_t i, bool desc,
                                         uint64_t& hi, uint32_t& lo) {
  lo = 0;
  if (KIND == KIND_I64) {
    hi = enc_i64(ld_stream(static_cast<const int64_t*>(col) + i), desc);
  } else if (KIND == KIND_U64) {
    hi = enc_u64(ld_stream(static_cast<const uint64_t*>(col) + i), desc);
  } else if (KIND == KIND_F64) {
    hi = enc_f64(ld_stream(static_cast<const double*>(col) + i), desc);
  } else {
    const uint8_t* p = static_cast<const uint8_t*>(col) + i * width;
    uint64_t h = 0;
    uint32_t l = 0;
    if ((width & 1) == 0) {
      // even widths: rows are 2-byte aligned -> half as many loads
      const uint16_t* q = reinterpret_cast<const uint16_t*>(p);
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        uint32_t v = (2 * b < (int)width) ? __ldg(q + b) : 0;  // little-endian pair
        h = (h << 16) | ((v & 0xFF) << 8) | (v >> 8);
      }
      if (width > 8) {
        uint32_t v = __ldg(q + 4);
        l = ((v & 0xFF) << 8) | (v >> 8);
      }
    } else {
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        uint32_t v = (b < (int)width) ? __ldg(p + b) : 0;
        h = (h << 8) | v;
      }
      if (width > 8) l = __ldg(p + 8);
    }
    if (desc) {
      // complement only the significant bytes so padding stays constant
      uint64_t mask = width >= 8 ? ~0ull : ~((1ull << (64 - 8 * width)) - 1);
      h = (~h) & mask;
      uint32_t lmask = width >= 10 ? 0xFFFFu : (width == 9 ? 0xFFu : 0u);
      l = (~l) & lmask;
    }
    hi = h;
    lo = l;
  }
}


// 80-bit comparison of encoded keys
__device__ __forceinline__ bool key_less(uint64_t ah, uint32_t al, uint64_t bh, uint32_t bl) {
  return ah < bh || (ah == bh && al < bl);
}

}  // namespace ds
