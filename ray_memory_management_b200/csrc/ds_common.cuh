// Shared helpers for the ds_shuffle CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "ds_shuffle.h"

namespace ds {

void set_error(const char* fmt, ...);

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over the caller's workspace.  With base == nullptr it only
// measures, which is how the size-query phase of every entry point works.
struct Workspace {
  char* base;
  size_t off;
  explicit Workspace(void* p) : base(static_cast<char*>(p)), off(0) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * sizeof(T);
    return p;
  }
  size_t bytes() const { return align_up(off, 256); }
};

#define DS_CHECK_ARG(cond, msg)                         \
  do {                                                  \
    if (!(cond)) {                                      \
      ds::set_error("%s: %s", __func__, msg);           \
      return DS_ERR_INVALID_ARG;                        \
    }                                                   \
  } while (0)

#define DS_CUDA(expr)                                                        \
  do {                                                                       \
    cudaError_t e__ = (expr);                                                \
    if (e__ != cudaSuccess) {                                                \
      ds::set_error("%s: %s failed: %s", __func__, #expr,                    \
                    cudaGetErrorString(e__));                                \
      return DS_ERR_CUDA;                                                    \
    }                                                                        \
  } while (0)

#define DS_LAUNCH_CHECK() DS_CUDA(cudaGetLastError())

// Resolves the two-phase workspace protocol after the layout has been laid out on
// `w`: returns true when the caller should return DS_OK right away (size query).
inline int ws_resolve(const Workspace& w, void* ws, size_t* ws_bytes, const char* fn,
                      ds_status* st) {
  *st = DS_OK;
  if (ws_bytes == nullptr) {
    set_error("%s: ws_bytes must not be NULL", fn);
    *st = DS_ERR_INVALID_ARG;
    return 1;
  }
  if (ws == nullptr) {
    *ws_bytes = w.bytes();
    return 1;
  }
  if (*ws_bytes < w.bytes()) {
    set_error("%s: workspace too small: %zu < %zu bytes", fn, *ws_bytes, w.bytes());
    *st = DS_ERR_WORKSPACE;
    return 1;
  }
  if ((reinterpret_cast<uintptr_t>(ws) & 255) != 0) {
    set_error("%s: workspace must be 256-byte aligned", fn);
    *st = DS_ERR_INVALID_ARG;
    return 1;
  }
  return 0;
}

void note_launch();  // counts kernel launches (ds_debug_launch_count)

// Optional per-launch timing for bench.py's roofline figures: CUDA events recorded on the
// launching stream around a launch, per kernel class, read back by ds_debug_collect_timing
// after the caller has synchronised.  Measurement aid: not thread safe, off by default.
enum TimedKernel { TIMED_PASS = 0, TIMED_GATHER = 1, TIMED_KINDS = 2 };
struct LaunchTimer {
  int kind;
  cudaStream_t s;
  cudaEvent_t a = nullptr, b = nullptr;
  double bytes;
  LaunchTimer(int kind, cudaStream_t s, double algorithmic_bytes);
  void stop();
};
int sm_count();  // SMs of the current device (cached per device)

__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t lanemask_lt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}

// streaming (read-once) loads / stores: keep L1 for the data we reuse
template <typename T>
__device__ __forceinline__ T ld_stream(const T* p) { return __ldcs(p); }
template <typename T>
__device__ __forceinline__ void st_stream(T* p, T v) { __stcs(p, v); }

// ---- radix-order key encodings ------------------------------------------------
// Every key type is mapped to an unsigned integer whose natural order is the sort
// order; descending sorts complement it (equal keys stay equal -> stable).
__device__ __forceinline__ uint64_t enc_i64(int64_t v, bool desc) {
  uint64_t k = static_cast<uint64_t>(v) ^ 0x8000000000000000ull;
  return desc ? ~k : k;
}
__device__ __forceinline__ int64_t dec_i64(uint64_t k, bool desc) {
  if (desc) k = ~k;
  return static_cast<int64_t>(k ^ 0x8000000000000000ull);
}
__device__ __forceinline__ uint64_t enc_u64(uint64_t v, bool desc) { return desc ? ~v : v; }
// float64: NaNs last in both directions, -0.0 == +0.0 (Arrow's comparison).
__device__ __forceinline__ uint64_t enc_f64(double d, bool desc) {
  if (d != d) return ~0ull;
  d = d + 0.0;  // folds -0.0 into +0.0 (IEEE: -0 + +0 = +0), identity otherwise
  uint64_t b = static_cast<uint64_t>(__double_as_longlong(d));
  uint64_t k = (b >> 63) ? ~b : (b | 0x8000000000000000ull);
  // non-NaN k lies in [0x000F..F (-inf), 0xFFF0..0 (+inf)], and so does ~k: both
  // stay below the NaN code ~0.
  return desc ? ~k : k;
}

}  // namespace ds
