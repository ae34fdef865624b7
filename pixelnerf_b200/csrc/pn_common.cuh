// Shared helpers for the pixelnerf_b200 CUDA kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <atomic>
#include <stdio.h>

#include "../../include/pixelnerf_b200.h"

namespace pn {

// Thread-local last error text, returned by pn_last_error().
void set_error(const char* fmt, ...);
// Bumps the process-wide kernel launch counter behind pn_launch_count().
void count_launch(int n = 1);

#define PN_CHECK_ARG(cond, ...)            \
  do {                                     \
    if (!(cond)) {                         \
      pn::set_error(__VA_ARGS__);          \
      return PN_ERR_INVALID;               \
    }                                      \
  } while (0)

#define PN_CUDA_OK(expr)                                                          \
  do {                                                                            \
    cudaError_t _e = (expr);                                                      \
    if (_e != cudaSuccess) {                                                      \
      pn::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return PN_ERR_CUDA;                                                         \
    }                                                                             \
  } while (0)

#define PN_LAUNCH_OK(name)                                                        \
  do {                                                                            \
    cudaError_t _e = cudaGetLastError();                                          \
    if (_e != cudaSuccess) {                                                      \
      pn::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));     \
      return PN_ERR_CUDA;                                                         \
    }                                                                             \
  } while (0)

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// Kernel attributes (the dynamic shared memory opt-in) are per DEVICE: a process that renders on cuda:1 after
// cuda:0 has to set them again there.  `seen` is one bit per device ordinal; setting an attribute twice is
// harmless, so the check-then-mark race between two threads of one device only repeats the call.
struct DeviceOnce {
  std::atomic<unsigned long long> seen{0};
  bool needed(int* dev_out) const {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { *dev_out = -1; return true; }
    *dev_out = dev;
    return !((seen.load(std::memory_order_acquire) >> dev) & 1ull);
  }
  void mark(int dev) {
    if (dev >= 0) seen.fetch_or(1ull << dev, std::memory_order_release);
  }
};

// ---- counter-based RNG (Philox4x32-10), used when the caller does not inject draws ----
struct Philox {
  uint32_t key[2];
  __device__ __forceinline__ Philox(uint64_t seed) {
    key[0] = (uint32_t)seed;
    key[1] = (uint32_t)(seed >> 32);
  }
  __device__ __forceinline__ uint4 operator()(uint64_t ctr_lo, uint64_t ctr_hi) const {
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32);
    uint32_t c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
    uint32_t k0 = key[0], k1 = key[1];
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      c0 = hi1 ^ c1 ^ k0;
      c1 = lo1;
      c2 = hi0 ^ c3 ^ k1;
      c3 = lo0;
      k0 += 0x9E3779B9u;
      k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }  // [0,1)

// Streams used to decorrelate the four draws of one render call.
enum { PN_STREAM_COARSE = 0, PN_STREAM_BIN = 1, PN_STREAM_IN = 2, PN_STREAM_NOISE = 3 };

__device__ __forceinline__ float rand_uniform(uint64_t seed, int stream, uint64_t idx) {
  uint4 r = Philox(seed)(idx >> 2, (uint64_t)stream);
  uint32_t v = (idx & 3) == 0 ? r.x : (idx & 3) == 1 ? r.y : (idx & 3) == 2 ? r.z : r.w;
  return u01(v);
}
__device__ __forceinline__ float rand_normal(uint64_t seed, int stream, uint64_t idx) {
  uint4 r = Philox(seed)(idx >> 1, (uint64_t)stream);
  float u1 = 1.0f - u01((idx & 1) ? r.z : r.x);  // (0,1]
  float u2 = u01((idx & 1) ? r.w : r.y);
  return sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
}

}  // namespace pn
