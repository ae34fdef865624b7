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

// ---- where a kernel's rays come from -------------------------------------------------------------------------
// Either an explicit (R, 8) tensor [origin, unit direction, near, far] (src/util/util.py:274-276), or -- SURVEY 8f.1,
// ray generation fused into the renderer -- a set of frames: ray r of the call is pixel first_ray + r of the
// concatenated (n_frames, H, W) pixel grid, and its origin / direction are computed where they are needed from the
// frame's camera-to-world pose and the intrinsics (util.unproj_map / gen_rays, src/util/util.py:113-143, 238-276), with
// exactly the arithmetic of gen_rays_kernel (csrc/pn_frame.cu), so both sources give bit-identical results.
struct RaySrc {
  const float* rays;      // non-null: explicit rays; everything below is unused
  const float* poses;     // (n_frames, pose_stride) row-major camera-to-world, 4 floats per row (3 or 4 rows)
  int pose_stride;
  int width, height;
  float fx, fy, cx, cy, z_near, z_far;
  long long first_ray;
};
static inline RaySrc ray_src(const float* rays) {
  RaySrc s = {};
  s.rays = rays;
  return s;
}
__device__ __forceinline__ float2 ray_near_far(const RaySrc& s, long long r) {
  if (s.rays) return *reinterpret_cast<const float2*>(s.rays + r * 8 + 6);
  return make_float2(s.z_near, s.z_far);
}
__device__ __forceinline__ float ray_far(const RaySrc& s, long long r) { return s.rays ? s.rays[r * 8 + 7] : s.z_far; }
// a = [ox, oy, oz, dx], b = [dy, dz, near, far]
__device__ __forceinline__ void ray_load(const RaySrc& s, long long r, float4& a, float4& b) {
  if (s.rays) {
    a = __ldg(reinterpret_cast<const float4*>(s.rays + r * 8));
    b = __ldg(reinterpret_cast<const float4*>(s.rays + r * 8 + 4));
    return;
  }
  const long long i = s.first_ray + r;
  const long long per = (long long)s.width * s.height;
  const int img = (int)(i / per);
  const int pix = (int)(i - (long long)img * per);
  const int py = pix / s.width, px = pix - py * s.width;
  // unproj_map: ((x - cx) / fx, -(y - cy) / fy, -1), normalised
  const float X = __fdiv_rn((float)px - s.cx, s.fx);
  const float Y = -__fdiv_rn((float)py - s.cy, s.fy);
  const float Z = -1.0f;
  const float n = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(X, X), __fmul_rn(Y, Y)), 1.0f));
  const float ux = __fdiv_rn(X, n), uy = __fdiv_rn(Y, n), uz = __fdiv_rn(Z, n);
  const float* P = s.poses + (size_t)img * s.pose_stride;
  // explicit intrinsics: one rounding sequence wherever this is inlined (no compiler-chosen FMA contraction)
  const float dx = __fmaf_rn(P[2], uz, __fmaf_rn(P[1], uy, __fmul_rn(P[0], ux)));
  const float dy = __fmaf_rn(P[6], uz, __fmaf_rn(P[5], uy, __fmul_rn(P[4], ux)));
  const float dz = __fmaf_rn(P[10], uz, __fmaf_rn(P[9], uy, __fmul_rn(P[8], ux)));
  a = make_float4(P[3], P[7], P[11], dx);
  b = make_float4(dy, dz, s.z_near, s.z_far);
}

// RaySrc forms of the entry points (defined beside their kernels; the extern "C" functions wrap them)
int sample_coarse_src(const RaySrc& rays, const float* u, const float* lin, float* z, int64_t n_rays, int n_coarse,
                      int lindisp, uint64_t seed, pn_stream_t stream);
int sample_fine_src(const RaySrc& rays, const float* z_coarse, const float* weights, const float* depth,
                    const float* u_bin, const float* u_in, const float* noise, float* z_sorted, float* z_imp,
                    float* z_depth, int64_t n_rays, int n_coarse, int n_imp, int n_depth, float depth_std, int lindisp,
                    uint64_t seed, pn_stream_t stream);
int point_taps_src(const pn_camera_t* cam, const RaySrc& rays, const float* z, int n_samples, const float* xyz,
                   const float* viewdirs, int64_t points_per_object, void* feat_x, bool grouped, pn_stream_t stream,
                   void* feat_x_lo = nullptr);
int composite_fwd_src(const float* rgb_sigma, const float* z, const RaySrc& rays, float* weights, float* rgb, float* depth,
                      int64_t n_rays, int n_samples, int white_bkgd, pn_stream_t stream);

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
