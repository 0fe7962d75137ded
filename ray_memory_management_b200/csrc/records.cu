// GraySort-format record files either side of the sort path.
//
// The reference only states the format (python/ray/experimental/raysort/constants.py:8:
// RECORD_SIZE = 100, 10-byte key first) and shells out to gensort / valsort binaries that are
// not in its tree (raysort/main.py:148-176,399-432).  These kernels move between the on-disk
// layout -- records of record_bytes back to back -- and the two Arrow columns the sort works on
// ({key: fixed_size_binary(key_bytes), payload: fixed_size_binary(record_bytes - key_bytes)}),
// and compute valsort's summary: records out of order, duplicate keys, and an
// order-independent checksum (sum of the CRC-32 of every record).
#include "ds_common.cuh"

namespace ds {

// One block moves 256 records through shared memory so that both sides use coalesced
// 4-byte accesses where alignment allows (record_bytes % 4 == 0 for GraySort's 100).
constexpr int REC_PER_BLOCK = 256;

__global__ void __launch_bounds__(256)
    split_records_kernel(const uint8_t* __restrict__ rec, uint32_t record_bytes, uint32_t key_bytes,
                         uint32_t n, uint8_t* __restrict__ key, uint8_t* __restrict__ payload) {
  extern __shared__ uint8_t sm[];
  const uint32_t r0 = blockIdx.x * REC_PER_BLOCK;
  const uint32_t rows = min(static_cast<uint32_t>(REC_PER_BLOCK), n - r0);
  const uint32_t bytes = rows * record_bytes;
  const uint8_t* src = rec + static_cast<uint64_t>(r0) * record_bytes;
  for (uint32_t o = threadIdx.x; o < bytes; o += blockDim.x) sm[o] = src[o];
  __syncthreads();
  const uint32_t pay = record_bytes - key_bytes;
  uint8_t* kd = key + static_cast<uint64_t>(r0) * key_bytes;
  for (uint32_t o = threadIdx.x; o < rows * key_bytes; o += blockDim.x) {
    const uint32_t r = o / key_bytes;
    kd[o] = sm[r * record_bytes + (o - r * key_bytes)];
  }
  uint8_t* pd = payload + static_cast<uint64_t>(r0) * pay;
  for (uint32_t o = threadIdx.x; o < rows * pay; o += blockDim.x) {
    const uint32_t r = o / pay;
    pd[o] = sm[r * record_bytes + key_bytes + (o - r * pay)];
  }
}

__global__ void __launch_bounds__(256)
    join_records_kernel(const uint8_t* __restrict__ key, const uint8_t* __restrict__ payload,
                        uint32_t record_bytes, uint32_t key_bytes, uint32_t n, uint8_t* __restrict__ rec) {
  extern __shared__ uint8_t sm[];
  const uint32_t r0 = blockIdx.x * REC_PER_BLOCK;
  const uint32_t rows = min(static_cast<uint32_t>(REC_PER_BLOCK), n - r0);
  const uint32_t pay = record_bytes - key_bytes;
  const uint8_t* ks = key + static_cast<uint64_t>(r0) * key_bytes;
  for (uint32_t o = threadIdx.x; o < rows * key_bytes; o += blockDim.x) {
    const uint32_t r = o / key_bytes;
    sm[r * record_bytes + (o - r * key_bytes)] = ks[o];
  }
  const uint8_t* ps = payload + static_cast<uint64_t>(r0) * pay;
  for (uint32_t o = threadIdx.x; o < rows * pay; o += blockDim.x) {
    const uint32_t r = o / pay;
    sm[r * record_bytes + key_bytes + (o - r * pay)] = ps[o];
  }
  __syncthreads();
  uint8_t* dst = rec + static_cast<uint64_t>(r0) * record_bytes;
  for (uint32_t o = threadIdx.x; o < rows * record_bytes; o += blockDim.x) dst[o] = sm[o];
}

__device__ __forceinline__ uint32_t crc32_byte(uint32_t crc, uint8_t b) {
  crc ^= b;
#pragma unroll
  for (int k = 0; k < 8; ++k) crc = (crc >> 1) ^ (0xEDB88320u & (0u - (crc & 1u)));
  return crc;
}

// out[0] += sum of CRC-32 over records, out[1] += records whose key is smaller than the
// previous record's key, out[2] += records whose key equals the previous record's key.
__global__ void __launch_bounds__(256)
    validate_records_kernel(const uint8_t* __restrict__ key, const uint8_t* __restrict__ payload,
                            uint32_t key_bytes, uint32_t pay_bytes, uint32_t n,
                            unsigned long long* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long sum = 0, unordered = 0, dup = 0;
  if (i < n) {
    uint32_t crc = 0xFFFFFFFFu;
    const uint8_t* k = key + static_cast<uint64_t>(i) * key_bytes;
    for (uint32_t b = 0; b < key_bytes; ++b) crc = crc32_byte(crc, k[b]);
    const uint8_t* p = payload + static_cast<uint64_t>(i) * pay_bytes;
    for (uint32_t b = 0; b < pay_bytes; ++b) crc = crc32_byte(crc, p[b]);
    sum = crc ^ 0xFFFFFFFFu;
    if (i > 0) {
      const uint8_t* q = k - key_bytes;
      int cmp = 0;
      for (uint32_t b = 0; b < key_bytes && cmp == 0; ++b) cmp = (int)k[b] - (int)q[b];
      unordered = cmp < 0;
      dup = cmp == 0;
    }
  }
  // block reduction, one atomic per counter per block
  __shared__ unsigned long long s[3][8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    unordered += __shfl_xor_sync(0xffffffffu, unordered, o);
    dup += __shfl_xor_sync(0xffffffffu, dup, o);
  }
  if (lane_id() == 0) {
    s[0][threadIdx.x >> 5] = sum;
    s[1][threadIdx.x >> 5] = unordered;
    s[2][threadIdx.x >> 5] = dup;
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    unsigned long long t = 0;
    for (int w = 0; w < 8; ++w) t += s[threadIdx.x][w];
    if (t) atomicAdd(&out[threadIdx.x], t);
  }
}

static ds_status check_records_args(const char* fn, uint32_t record_bytes, uint32_t key_bytes, uint64_t n) {
  if (n > 0xFFFFFFFFull) {
    set_error("%s: n exceeds 2^32-1", fn);
    return DS_ERR_TOO_LARGE;
  }
  if (key_bytes == 0 || key_bytes >= record_bytes || record_bytes > 192) {
    set_error("%s: need 0 < key_bytes < record_bytes <= 192 (got %u, %u)", fn, key_bytes, record_bytes);
    return DS_ERR_INVALID_ARG;
  }
  return DS_OK;
}

}  // namespace ds

extern "C" {

ds_status ds_split_records(const uint8_t* records, uint32_t record_bytes, uint32_t key_bytes, uint64_t n,
                           uint8_t* key_out, uint8_t* payload_out, void* stream) {
  using namespace ds;
  ds_status st = check_records_args(__func__, record_bytes, key_bytes, n);
  if (st != DS_OK || n == 0) return st;
  if (!records || !key_out || !payload_out) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  const uint32_t blocks = (static_cast<uint32_t>(n) + REC_PER_BLOCK - 1) / REC_PER_BLOCK;
  note_launch();
  split_records_kernel<<<blocks, 256, REC_PER_BLOCK * record_bytes, static_cast<cudaStream_t>(stream)>>>(
      records, record_bytes, key_bytes, static_cast<uint32_t>(n), key_out, payload_out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_join_records(const uint8_t* key, const uint8_t* payload, uint32_t record_bytes,
                          uint32_t key_bytes, uint64_t n, uint8_t* records_out, void* stream) {
  using namespace ds;
  ds_status st = check_records_args(__func__, record_bytes, key_bytes, n);
  if (st != DS_OK || n == 0) return st;
  if (!key || !payload || !records_out) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  const uint32_t blocks = (static_cast<uint32_t>(n) + REC_PER_BLOCK - 1) / REC_PER_BLOCK;
  note_launch();
  join_records_kernel<<<blocks, 256, REC_PER_BLOCK * record_bytes, static_cast<cudaStream_t>(stream)>>>(
      key, payload, record_bytes, key_bytes, static_cast<uint32_t>(n), records_out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_validate_records(const uint8_t* key, const uint8_t* payload, uint32_t record_bytes,
                              uint32_t key_bytes, uint64_t n, uint64_t* out3, void* stream) {
  using namespace ds;
  ds_status st = check_records_args(__func__, record_bytes, key_bytes, n);
  if (st != DS_OK || n == 0) return st;
  if (!key || !payload || !out3) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  note_launch();
  validate_records_kernel<<<(static_cast<uint32_t>(n) + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      key, payload, key_bytes, record_bytes - key_bytes, static_cast<uint32_t>(n),
      reinterpret_cast<unsigned long long*>(out3));
  DS_LAUNCH_CHECK();
  return DS_OK;
}

}  // extern "C"
