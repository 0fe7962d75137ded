// Library-wide pieces: version, thread-local error text, device properties, and the
// splitter-sampling kernel.
#include <stdarg.h>

#include <atomic>
#include <vector>

#include "ds_common.cuh"
#include "ds_keys.cuh"

namespace ds {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static std::atomic<unsigned long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
unsigned long long launch_count() { return g_launches.load(std::memory_order_relaxed); }

static bool g_timing = false;
struct TimedLaunch {
  cudaEvent_t a, b;
  double bytes;
};
static std::vector<TimedLaunch> g_timed[TIMED_KINDS];

LaunchTimer::LaunchTimer(int k, cudaStream_t st, double algorithmic_bytes) : kind(k), s(st), bytes(algorithmic_bytes) {
  if (!g_timing) return;
  if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) {
    a = b = nullptr;
    return;
  }
  cudaEventRecord(a, s);
}
void LaunchTimer::stop() {
  if (!a) return;
  cudaEventRecord(b, s);
  g_timed[kind].push_back(TimedLaunch{a, b, bytes});
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  int& c = cached[dev & 63];
  if (c == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    c = v;
  }
  return c;
}

// One warp per sample: sample j copies one row chosen inside stratum j.
__global__ void sample_rows_kernel(const uint8_t* __restrict__ col, uint32_t row_bytes, uint64_t n,
                                   uint32_t k, uint64_t seed, uint8_t* __restrict__ out) {
  const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (j >= k) return;
  const uint64_t lo = static_cast<uint64_t>((static_cast<unsigned __int128>(j) * n) / k);
  const uint64_t hi = static_cast<uint64_t>((static_cast<unsigned __int128>(j + 1) * n) / k);
  const uint64_t row = lo + philox64(seed, 0x73616d70ull /* "samp" */, j) % (hi - lo);
  const uint8_t* src = col + row * row_bytes;
  uint8_t* dst = out + static_cast<uint64_t>(j) * row_bytes;
  for (uint32_t b = lane_id(); b < row_bytes; b += 32) dst[b] = src[b];
}

}  // namespace ds

extern "C" {

int32_t ds_version(void) { return 100; /* 0.1.0 */ }

const char* ds_last_error(void) { return ds::g_err; }

// Per-launch timing of a kernel class (0 = radix pass, 1 = row gather), see LaunchTimer.
void ds_debug_pass_timing(int32_t enable) { ds::g_timing = enable != 0; }

// Sums duration (ms) and algorithmic bytes of the recorded launches of `kind` and clears the
// record.  The caller must have synchronised the stream(s) first.
int32_t ds_debug_collect_timing(int32_t kind, double* total_ms, int32_t* launches, double* bytes) {
  double tot = 0, by = 0;
  int cnt = 0;
  if (kind < 0 || kind >= ds::TIMED_KINDS) return -1;
  for (auto& ev : ds::g_timed[kind]) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, ev.a, ev.b) == cudaSuccess) {
      tot += ms;
      by += ev.bytes;
      ++cnt;
    }
    cudaEventDestroy(ev.a);
    cudaEventDestroy(ev.b);
  }
  ds::g_timed[kind].clear();
  if (total_ms) *total_ms = tot;
  if (launches) *launches = cnt;
  if (bytes) *bytes = by;
  return 0;
}

// Number of kernels this library has launched in this process (bench.py: gpu_launches).
uint64_t ds_debug_launch_count(void) { return ds::launch_count(); }

ds_status ds_sample_rows(const void* col, uint32_t row_bytes, uint64_t n, uint32_t k, uint64_t seed,
                         void* out, void* stream) {
  if (k == 0) return DS_OK;
  if (!col || !out || row_bytes == 0) {
    ds::set_error("%s: bad argument", __func__);
    return DS_ERR_INVALID_ARG;
  }
  if (k > n) {
    ds::set_error("%s: k=%u exceeds n=%llu (sampling is without replacement)", __func__, k,
                  (unsigned long long)n);
    return DS_ERR_INVALID_ARG;
  }
  const uint32_t threads = 128;
  const uint32_t blocks = (k * 32u + threads - 1) / threads;
  ds::note_launch(); ds::sample_rows_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(col), row_bytes, n, k, seed, static_cast<uint8_t*>(out));
  DS_LAUNCH_CHECK();
  return DS_OK;
}

}  // extern "C"
